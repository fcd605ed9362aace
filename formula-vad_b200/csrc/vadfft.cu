// vadfft.cu -- K3b: 1:D linear upsample of the denoised stream, re-blocking into non-overlapping
// 1024-sample frames, periodic-Hann real FFT-1024, magnitude * 1/sum(w), speech-band sum, min over
// channels; plus the sample-weighted raw inter-channel volume ratio of the frame.
//
// Replaces (reference):
//   resample.upsampleAudio / interpolate  (audio_utils/resample.zig:32-79; lerp is a fused a+(b-a)*t)
//   BufferedFFT.write re-blocking         (AudioPipeline/BufferedFFT.zig:129-160, SegmentWriter.zig:46-109)
//   BufferedFFT.fft                       (BufferedFFT.zig:162-181) -> FFT.fft -> kiss_fftr 1024
//   BufferedFFT.averageVolumeInBand       (BufferedFFT.zig:183-202)  sequential f32 sum of bins lo..hi
//   min over channels                     (AudioPipeline/VADMachine.zig:153-158, start value 999)
//   BufferedVolumeAnalyzer.analyseVolume  (BufferedVolumeAnalyzer.zig:48-69) from the chunk sums of squares
//   VADMetadata weighted merge            (VADMetadata.zig:29-60) through the three buffered steps
//
// One CTA per (VAD frame, stream); warp w handles channel w (channels beyond the warp count loop).
#include <algorithm>

#include "fft_dev.cuh"
#include "kernels.cuh"

namespace fvad {

constexpr int kVadWarps = 8;
constexpr int kVadThreads = 32 * kVadWarps;
constexpr int kMaxBands = 8;

// ratio of one chunk as the reference sees it after the analyser and the denoiser step have each
// pushed it with weight `chunk` and divided it out again (f32, no contraction)
__device__ __forceinline__ float chunk_ratio(const float *__restrict__ sumsq, int C, int s, int pitch, int slot, int chunk) {
    float vmin = 1.0f, vmax = 0.0f;
    for (int ch = 0; ch < C; ++ch) {
        const float ss = sumsq[(size_t)(s * C + ch) * pitch + slot];
        const float v = sqrtf(__fdiv_rn(ss, (float)chunk));  // audio_utils.zig:22-23
        if (v < vmin) vmin = v;
        if (v > vmax) vmax = v;
    }
    float r = (vmax == 0.0f) ? 0.0f : __fdiv_rn(vmin, vmax);
    const float w = (float)chunk;
    r = __fdiv_rn(__fmul_rn(r, w), w);  // BufferedVolumeAnalyzer.write :35-43
    r = __fdiv_rn(__fmul_rn(r, w), w);  // BufferedDenoiser.write :83-86,114
    return r;
}

// One warp per (VAD frame, stream) item, eight warps per CTA, CTAs loop over items: the 8 KB twiddle
// table is staged once per CTA and every lane-constant operand (FFT twiddles, Hann weights) sits in
// registers across items.  The band sum is a per-lane partial + xor reduction (the reference sums the
// 33 bins sequentially, BufferedFFT.zig:190-199; the difference is f32 summation order, ~1e-7 relative,
// below the FFT's own rounding distance to kissfft).
// rfft_bin over the padded Z (element k lives at k + k/16)
__device__ __forceinline__ float2 rfft_bin_padded(const float2 *__restrict__ Z, const float2 *__restrict__ T, int N, int k) {
    const int ia = (k == N) ? 0 : k, ib = (k == 0) ? 0 : N - k;
    const float2 a = Z[ia + (ia >> 4)];
    const float2 bb = Z[ib + (ib >> 4)];
    const float2 b = make_float2(bb.x, -bb.y);
    const float2 e = cadd(a, b), d = csub(a, b);
    const float2 t = T[k];
    const float2 o = cmul(d, make_float2(t.y, -t.x));
    return make_float2((e.x + o.x) * 0.5f, (k == 0 || k == N) ? 0.0f : (e.y + o.y) * 0.5f);
}

template <int D>
__global__ void __launch_bounds__(kVadThreads, 2) vadfft_kernel(VadFftArgs a, const float2 *__restrict__ gT,
                                                                const float *__restrict__ ghann, float norm) {
    __shared__ float2 T[1024];
    __shared__ float2 zb[kVadWarps][512 + 32];  // one pad slot per 16: the bit-reversed store (stride 16) spreads over all banks
    __shared__ float2 hann2[512];

    const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // (tells the compiler the warp index is uniform)
    for (int i = tid; i < 1024; i += kVadThreads) T[i] = gT[i];
    for (int i = tid; i < 512; i += kVadThreads) hann2[i] = reinterpret_cast<const float2 *>(ghann)[i];
    __syncthreads();

    WarpFftTw<16, false> ftw;
    ftw.init(T, lane);

    const size_t ypitch = kYLook + (size_t)kChunk16 * a.K;
    const long long n_items = (long long)a.n_frames * a.S;
    float2 *Z = zb[warp];

    // Work queue: every warp starts on its own item, then draws kGrab more at a time from a counter.  (The VAD scan
    // of the previous span shares some SMs with this launch; with static striding those SMs' warps set the tail.)
    constexpr int kGrab = 4;
    const long long n_static = (long long)gridDim.x * kVadWarps;
    long long item = (long long)blockIdx.x * kVadWarps + warp, grab_end = item + 1;
    for (;; ++item) {
        if (item >= grab_end) {
            if (!a.work_ctr) { item += n_static - 1; grab_end = item + 1; }
            else {
                unsigned long long g = 0;
                if (lane == 0) g = atomicAdd(a.work_ctr, (unsigned long long)kGrab);
                item = n_static + (long long)__shfl_sync(0xffffffffu, g, 0);
                grab_end = item + kGrab;
            }
        }
        if (item >= n_items) break;
        const int fi = (int)(item % a.n_frames);  // frame within this launch (fastest: neighbouring warps read neighbouring audio)
        const int s = (int)(item / a.n_frames);
        const long long frame = a.first_frame + fi;
        const long long s0 = frame * 1024;                 // absolute start sample (input rate)
        const long long rel = s0 - a.span_start;           // relative to the span start, may be negative

        float bmin[kMaxBands];
#pragma unroll
        for (int b = 0; b < kMaxBands; ++b) bmin[b] = 999.0f;  // VADMachine.zig:153

        for (int ch = 0; ch < a.C; ++ch) {
            // y16 index 0 of the span sits at +kYLook; (rel + i) is the input-rate offset
            const float *__restrict__ y = a.y16 + (size_t)(s * a.C + ch) * ypitch + kYLook;
            float2 v[16];
            if (D == 1) {
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const int n = lane + 32 * q;
                    const float2 t = *reinterpret_cast<const float2 *>(y + rel + 2 * n);  // rel and the row pitch are even
                    const float2 hw = hann2[n];
                    v[q] = make_float2(t.x * hw.x, t.y * hw.y);
                }
            } else {
                // resample.zig:44-59: input-rate position 3m+2 carries y16[m]; 3m and 3m+1 interpolate from y16[m-1] with
                // frac = (phase+1)/3 in f32.  The frame's 1024 positions draw on y16[m0-1 .. m0+341]: the warp stages that
                // run in its (not yet used) Z buffer with coalesced loads, then every lane interpolates its 32 positions
                // from shared memory -- no 64-bit division, no per-sample global loads, no per-sample f32 division.
                const int rel32 = (int)rel;                                                  // |rel| < 2^31 (a span is < 2^27 samples)
                const int m0 = rel32 >= 0 ? rel32 / D : -((-rel32 + D - 1) / D);            // floor(rel / D)
                const int r0 = rel32 - m0 * D;                                               // phase of the frame's first position
                float *ys = reinterpret_cast<float *>(Z);
                const int m_hi = kChunk16 * a.K - 1;                                         // last y16 sample of the span
#pragma unroll
                for (int i = 0; i < 11; ++i) {
                    const int m = m0 - 1 + lane + 32 * i;
                    ys[lane + 32 * i] = m <= m_hi ? y[m] : 0.0f;
                }
                __syncwarp();
                const float f1 = __fdiv_rn(1.0f, (float)D), f2 = __fdiv_rn(2.0f, (float)D);
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const int n = lane + 32 * q;
                    float e[2];
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const int u = r0 + 2 * n + j;      // position relative to 3*m0
                        const int dm = u / D, ph = u - dm * D;
                        const float prev = ys[dm], cur = ys[dm + 1];  // y16[m0+dm-1], y16[m0+dm]
                        e[j] = ph == D - 1 ? cur : fmaf(cur - prev, ph == 0 ? f1 : f2, prev);
                    }
                    const float2 hw = hann2[n];
                    v[q] = make_float2(e[0] * hw.x, e[1] * hw.y);
                }
                __syncwarp();  // the staged run is read before the transform's result overwrites it
            }
            warp_cfft_tw<16, false>(v, ftw);
#pragma unroll
            for (int k1 = 0; k1 < 16; ++k1) Z[k1 + 17 * bitrev5(lane)] = v[k1];  // = pad(k1 + 16*bitrev5(lane))
            __syncwarp();
            for (int b = 0; b < a.n_bands; ++b) {
                const int lo = a.band_lo[b], hi = a.band_hi[b];
                float acc = 0.0f;
                for (int k = lo + lane; k <= hi; k += 32) {
                    const float2 X = rfft_bin_padded(Z, T, 512, k);
                    // FFT.zig:16-18 abs = sqrt(r*r + i*i); BufferedFFT.zig:176 * norm_factor
                    acc += __fmul_rn(sqrtf(__fadd_rn(__fmul_rn(X.x, X.x), __fmul_rn(X.y, X.y))), norm);
                }
#pragma unroll
                for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
                bmin[b] = fminf(bmin[b], acc);
            }
            __syncwarp();
        }
        if (lane == 0) {
            for (int b = 0; b < a.n_bands; ++b) a.band[((size_t)b * a.S + s) * a.max_frames + fi] = bmin[b];
            // VADMetadata merge over the (at most two) chunks this frame draws samples from
            const int pitch = 1 + a.K;
            const long long e0 = s0 + 1023;
            const long long c_first = (s0 >= a.span_start) ? (rel / a.chunk) : -1;     // -1 = previous span's last chunk
            const long long c_last = (e0 - a.span_start) / a.chunk;                    // e0 >= span_start always
            float sum = 0.0f, wsum = 0.0f;
            if (c_first == c_last) {
                const float r = chunk_ratio(a.sumsq, a.C, s, pitch, (int)c_first + 1, a.chunk);
                sum = __fmul_rn(r, 1024.0f);
                wsum = 1024.0f;
            } else {
                const long long boundary = a.span_start + c_last * a.chunk;  // first sample of the later chunk
                const float n0 = (float)(boundary - s0), n1 = (float)(s0 + 1024 - boundary);
                const float r0 = chunk_ratio(a.sumsq, a.C, s, pitch, (int)c_first + 1, a.chunk);
                const float r1 = chunk_ratio(a.sumsq, a.C, s, pitch, (int)c_last + 1, a.chunk);
                sum = __fadd_rn(__fmul_rn(r0, n0), __fmul_rn(r1, n1));
                wsum = __fadd_rn(n0, n1);
            }
            a.ratio[(size_t)s * a.max_frames + fi] = __fdiv_rn(sum, wsum);
        }
    }
}

void launch_vadfft(const VadFftArgs &a, const Tables &t, cudaStream_t s) {
    if (a.n_frames <= 0 || a.S <= 0) return;
    if (a.n_bands > kMaxBands) throw Error(FVAD_ERR_INVALID_ARGUMENT, "too many distinct speech bands");
    if (a.fft_size != 1024) throw Error(FVAD_ERR_INVALID_FFT_SIZE, "GPU VAD path supports fft_size 1024");
    const long long items = (long long)a.n_frames * a.S;
    const unsigned grid = (unsigned)std::min<long long>((items + kVadWarps - 1) / kVadWarps, 148 * 2);
    if (a.work_ctr) FVAD_CUDA_CHECK(cudaMemsetAsync(a.work_ctr, 0, sizeof(unsigned long long), s));
    if (a.D == 1) vadfft_kernel<1><<<grid, kVadThreads, 0, s>>>(a, t.T1024, t.hann1024, t.vad_norm);
    else if (a.D == 3) vadfft_kernel<3><<<grid, kVadThreads, 0, s>>>(a, t.T1024, t.hann1024, t.vad_norm);
    else throw Error(FVAD_ERR_INVALID_SAMPLE_RATE, "unsupported upsample factor");
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace fvad
