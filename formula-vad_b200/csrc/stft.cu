// stft.cu -- K1: decimate + framing + sqrt-Hann window + real FFT-320 + log-power features,
// fused with the per-chunk sum of squares of the raw input (the volume analyser's RMS).
//
// Replaces, per chunk of one channel (reference src/NSNet2.zig:196-219):
//   resample.downsampleAudio      (audio_utils/resample.zig:9-29)   pick every D-th sample
//   NSNet2.calcSpectrogram        (NSNet2.zig:239-264) -> FFT.fft (FFT.zig:85-113) -> kiss_fftr
//   NSNet2.calcFeatures           (NSNet2.zig:266-287) log10(max(re^2+im^2, 1e-12))
//   audio_utils.rmsVolume         (audio_utils.zig:14-24) -- only the sum of squares here
//
// One CTA per (chunk, stream-channel): the chunk's D*8000 input samples are read from HBM exactly
// once with 128-bit loads, squared-and-summed, decimated into shared memory (51 hops = 160-sample
// halo + 8000), then ten warps run five 320-point frames each through the register/shuffle FFT.
// Algorithmic HBM bytes per frame: 160*D*4 read + 161*4 features (+ 161*8 spectrum) written.
#include "fft_dev.cuh"
#include "kernels.cuh"

namespace fvad {

constexpr int kStftThreads = 320;
constexpr int kStftWarps = kStftThreads / 32;

template <int D>
__global__ void __launch_bounds__(kStftThreads) stft_kernel(StftArgs a, const float2 *__restrict__ gT,
                                                            const float *__restrict__ gwin) {
    __shared__ __align__(16) float xs[(kFrames + 1) * kHop];  // decimated: halo hop + 50 hops
    __shared__ float2 T[256];  // indices used by the N=160 transform stay below 249
    __shared__ float win[kNfft];
    __shared__ float2 zb[kStftWarps][kHop];
    __shared__ float red[kStftWarps];

    const int c = blockIdx.x;    // chunk within the span
    const int sc = blockIdx.y;   // stream-channel
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    for (int i = tid; i < kNfft; i += kStftThreads) {
        if (i < 256) T[i] = gT[i];
        win[i] = gwin[i];
    }

    // ---- load + decimate + sum of squares ----
    const float *__restrict__ x = a.x[sc] + a.x_offset + (size_t)c * (kChunk16 * D);
    float ss = 0.0f;
    {
        const float4 *x4 = reinterpret_cast<const float4 *>(x);
        constexpr int n4 = kChunk16 * D / 4;
        for (int i = tid; i < n4; i += kStftThreads) {
            const float4 v = __ldg(x4 + i);
            ss = fmaf(v.x, v.x, ss);
            ss = fmaf(v.y, v.y, ss);
            ss = fmaf(v.z, v.z, ss);
            ss = fmaf(v.w, v.w, ss);
            if (D == 1) {
                *reinterpret_cast<float4 *>(&xs[kHop + 4 * i]) = v;
            } else {
                // sample index s = 4i+j is kept when s % D == 0, at xs[hop + s/D]
                const int s0 = 4 * i;
                const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if ((s0 + j) % D == 0) xs[kHop + (s0 + j) / D] = e[j];
            }
        }
    }
    // halo: the 160 decimated samples that precede this chunk
    if (tid < kHop) {
        float h;
        if (c == 0) h = a.halo_in[(size_t)sc * kHop + tid];
        else h = __ldg(x - (size_t)(kHop - tid) * D);
        xs[tid] = h;
    }
    // block reduction of the sum of squares
#pragma unroll
    for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (lane == 0) red[warp] = ss;
    __syncthreads();
    if (tid == 0) {
        float t = 0.0f;
#pragma unroll
        for (int w = 0; w < kStftWarps; ++w) t += red[w];
        a.sumsq[(size_t)sc * (a.K + 1) + 1 + c] = t;  // slot 0 of each row = last chunk of the previous span
    }
    // the last chunk of the span leaves its last hop as the next span's halo
    if (c == a.K - 1 && tid < kHop) a.halo_out[(size_t)sc * kHop + tid] = xs[kFrames * kHop + tid];

    // ---- frames ----
    // lane-constant operands live in registers for all five frames of this warp
    WarpFftTw<5, false> ftw;
    ftw.init(T, lane);
    float2 wreg[5], st[3];
#pragma unroll
    for (int q = 0; q < 5; ++q) wreg[q] = *reinterpret_cast<const float2 *>(win + 2 * (lane + 32 * q));
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const float2 t = T[(lane + 32 * q) & 255];  // pairs k = lane + 32q <= 80 are used
        st[q] = make_float2(t.y, -t.x);             // super twiddle -i*T[k] (kiss_fftr.c:104)
    }
    float2 *Z = zb[warp];
    for (int f = warp; f < kFrames; f += kStftWarps) {
        const float *fx = xs + f * kHop;
        float2 v[5];
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            const float2 s2 = *reinterpret_cast<const float2 *>(fx + 2 * (lane + 32 * q));
            v[q] = make_float2(s2.x * wreg[q].x, s2.y * wreg[q].y);  // FFT.zig:188-196 window multiply
        }
        warp_cfft_tw<5, false>(v, ftw);
#pragma unroll
        for (int k1 = 0; k1 < 5; ++k1) Z[cfft_out_index<5>(lane, k1)] = v[k1];
        __syncwarp();
        const size_t row = (size_t)sc * (kFrames * a.K) + (size_t)c * kFrames + f;
        const size_t frow = (size_t)sc * (kCtx + kFrames * a.K) + kCtx + (size_t)c * kFrames + f;
        float2 *srow = a.spec ? reinterpret_cast<float2 *>(a.spec) + row * kBins : nullptr;
        float *frw = a.feat + frow * kBins;
        // NSNet2.zig:275-284: log10(max(re^2 + im^2, 1e-12)), squares and sum rounded separately (no FMA);
        // the logarithm is the hardware lg2 (abs error ~1e-7 in log10 units, far inside every parity gate)
        auto emit = [&](int k, float2 X) {
            if (srow) srow[k] = X;
            const float pw = __fadd_rn(__fmul_rn(X.x, X.x), __fmul_rn(X.y, X.y));
            frw[k] = __log2f(fmaxf(pw, 1e-12f)) * 0.30102999566398120f;
        };
        // real-FFT unpack in kissfft's pair form: lane handles bins k and 160-k, k = lane + 32q in [0, 80]
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const int k = lane + 32 * q;
            if (k <= kHop / 2) {
                const float2 zk = Z[k];
                if (k == 0) {  // DC and Nyquist (kiss_fftr.c:82-89)
                    emit(0, make_float2(zk.x + zk.y, 0.0f));
                    emit(kHop, make_float2(zk.x - zk.y, 0.0f));
                } else {
                    float2 Xk, Xnk;
                    rfft_pair(zk, Z[kHop - k], st[q], Xk, Xnk);
                    if (k != kHop - k) emit(k, Xk);  // at k = 80 kissfft's second store wins
                    emit(kHop - k, Xnk);
                }
            }
        }
        __syncwarp();
    }
}

void launch_stft(const StftArgs &a, const Tables &t, cudaStream_t s) {
    if (a.K <= 0 || a.SC <= 0) return;
    dim3 grid((unsigned)a.K, (unsigned)a.SC);
    if (a.D == 1) stft_kernel<1><<<grid, kStftThreads, 0, s>>>(a, t.T320, t.win320);
    else if (a.D == 3) stft_kernel<3><<<grid, kStftThreads, 0, s>>>(a, t.T320, t.win320);
    else throw Error(FVAD_ERR_INVALID_SAMPLE_RATE, "unsupported decimation factor");
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace fvad
