// stft.cu -- K1: decimate + framing + sqrt-Hann window + real FFT-320 + log-power features,
// fused with the per-chunk sum of squares of the raw input (the volume analyser's RMS).
//
// Replaces, per chunk of one channel (reference src/NSNet2.zig:196-219):
//   resample.downsampleAudio      (audio_utils/resample.zig:9-29)   pick every D-th sample
//   NSNet2.calcSpectrogram        (NSNet2.zig:239-264) -> FFT.fft (FFT.zig:85-113) -> kiss_fftr
//   NSNet2.calcFeatures           (NSNet2.zig:266-287) log10(max(re^2+im^2, 1e-12))
//   audio_utils.rmsVolume         (audio_utils.zig:14-24) -- only the sum of squares here
//
// A CTA works on one chunk at a time: the chunk's D*8000 input samples are read from HBM exactly
// once, squared-and-summed, decimated into shared memory (51 hops = 160-sample halo + 8000), then ten
// warps run five 320-point frames each through the register/shuffle FFT.  At D = 1 (16 kHz input) a
// CTA walks kChunksPerCta consecutive chunks and one elected thread prefetches the next chunk into
// the other half of a double-buffered tile with cp.async.bulk + mbarrier while the warps transform
// the current one, so no warp ever waits on HBM; at D = 3 the decimating loads are batched.
// Algorithmic HBM bytes per frame: 160*D*4 read + 161*4 features (+ 161*8 spectrum) written.
#include <cstdlib>

#include "fft_dev.cuh"
#include "kernels.cuh"
#include "tc_dev.cuh"

namespace fvad {

constexpr int kStftThreads = 320;
constexpr int kStftWarps = kStftThreads / 32;
constexpr int kXs = (kFrames + 1) * kHop;  // decimated tile: halo hop + 50 hops
constexpr int kChunksPerCta = 8;

// lane-constant operands of the frame transform, held in registers across frames
struct StftLaneConst {
    WarpFftTw<5, false> ftw;
    float2 wreg[5], st[3];
    __device__ __forceinline__ void init(const float2 *__restrict__ T, const float *__restrict__ win, int lane) {
        ftw.init(T, lane);
#pragma unroll
        for (int q = 0; q < 5; ++q) {  // window * 1/2: the real-FFT unpack's HALF_OF (kiss_fftr.c:111-114), applied up front
            const float2 w = *reinterpret_cast<const float2 *>(win + 2 * (lane + 32 * q));
            wreg[q] = make_float2(0.5f * w.x, 0.5f * w.y);
        }
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const float2 t = T[(lane + 32 * q) & 255];  // pairs k = lane + 32q <= 80 are used
            st[q] = make_float2(t.y, -t.x);             // super twiddle -i*T[k] (kiss_fftr.c:104)
        }
    }
};

__device__ __forceinline__ float lg2_ftz(float x) {
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// one windowed frame v (lane l holds samples 2(l+32q), 2(l+32q)+1) -> its spectrum row / feature rows
__device__ __forceinline__ void stft_emit(const StftArgs &a, const StftLaneConst &lc, float2 (&v)[5], float2 *__restrict__ Z,
                                          __half *__restrict__ fh, int sc, int c, int f, int lane) {
    warp_cfft_tw<5, false>(v, lc.ftw);
#pragma unroll
    for (int k1 = 0; k1 < 5; ++k1) Z[cfft_out_index<5>(lane, k1)] = v[k1];
    __syncwarp();
    const size_t row = (size_t)sc * (kFrames * a.K) + (size_t)c * kFrames + f;
    const size_t frow = (size_t)sc * (kCtx + kFrames * a.K) + kCtx + (size_t)c * kFrames + f;
    float2 *srow = a.spec ? reinterpret_cast<float2 *>(a.spec) + row * kBins : nullptr;
    float *frw = a.feat ? a.feat + frow * kBins : nullptr;
    __half *h16 = a.feat16 ? static_cast<__half *>(a.feat16) + frow * 176 : nullptr;
    // real-FFT unpack in kissfft's pair form (kiss_fftr.c:93-116): the lane handles bins k and 160-k for
    // k = lane + 32q in [0, 80].  k = 0 runs through the same arithmetic with Z[0] as its own partner
    // (gives DC = x + y, Nyquist = x - y, imaginary parts 0, kiss_fftr.c:82-89); at k = 80 both stores hit
    // bin 80 and the second wins, as in kissfft.
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const int k = lane + 32 * q;
        if (q < 2 || k <= kHop / 2) {
            float2 X[2];
            rfft_pair_halved(Z[k], Z[k ? kHop - k : 0], lc.st[q], X[0], X[1]);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int kk = h ? kHop - k : k;
                if (srow) srow[kk] = X[h];
                // NSNet2.zig:275-284: log10(max(re^2 + im^2, 1e-12)), squares and sum rounded separately (no
                // FMA); the logarithm is the hardware lg2 (abs error ~1e-7 in log10 units, inside every gate).  The
                // operand is >= 1e-12, never subnormal, so the .ftz form (one MUFU, no rescale sequence) is exact here.
                const float2 sq = cscale(X[h], X[h]);  // both squares in one packed multiply, each rounded on its own
                const float pw = __fadd_rn(sq.x, sq.y);
                const float ft = lg2_ftz(fmaxf(pw, 1e-12f)) * 0.30102999566398120f;
                if (frw) frw[kk] = ft;
                if (h16) fh[kk] = __float2half_rn(ft);
            }
        }
    }
    __syncwarp();
    if (h16 && lane < 22)  // the row leaves as 22 consecutive 16-byte stores
        *reinterpret_cast<uint4 *>(h16 + 8 * lane) = *reinterpret_cast<const uint4 *>(fh + 8 * lane);
    __syncwarp();
}

// frames f = warp, warp+10, ... of the tile xs -> spectrum rows / feature rows
__device__ __forceinline__ void stft_frames(const StftArgs &a, const StftLaneConst &lc, const float *__restrict__ xs,
                                            float2 *__restrict__ Z, __half *__restrict__ fh, int sc, int c, int warp, int lane) {
    // fh: this warp's 176-half staging row for the fp16 feature rows (bins 161..175 stay zero)
    if (a.feat16 && lane < 15) fh[kBins + lane] = __float2half(0.0f);
    for (int f = warp; f < kFrames; f += kStftWarps) {
        const float *fx = xs + f * kHop;
        float2 v[5];
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            const float2 s2 = *reinterpret_cast<const float2 *>(fx + 2 * (lane + 32 * q));
            v[q] = cscale(s2, lc.wreg[q]);  // FFT.zig:188-196 window multiply (and the 1/2)
        }
        stft_emit(a, lc, v, Z, fh, sc, c, f, lane);
    }
}

// block-wide sum of the per-thread partial sums of squares -> sumsq slot of chunk c (thread 0 writes)
__device__ __forceinline__ void stft_sumsq(const StftArgs &a, float ss, float *red, int sc, int c, int tid) {
    const int lane = tid & 31, warp = tid >> 5;
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
}

// ---- D = 1: persistent over kChunksPerCta chunks, double-buffered bulk prefetch ----
struct StftSmem1 {
    float xs[2][kXs];          // 2 x 32 640 B
    float2 T[256];
    float win[kNfft];
    float2 zb[kStftWarps][kHop];
    float red[2][kStftWarps];
    __align__(16) __half fh[kStftWarps][176];
    uint64_t full[2];
};

__global__ void __launch_bounds__(kStftThreads, 2) stft_kernel_d1(StftArgs a, const float2 *__restrict__ gT,
                                                                  const float *__restrict__ gwin) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    StftSmem1 &sm = *reinterpret_cast<StftSmem1 *>(smem_raw);
    const int sc = blockIdx.y;
    const int c_begin = blockIdx.x * kChunksPerCta, c_end = min(a.K, c_begin + kChunksPerCta);
    const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // (tells the compiler the warp index is uniform)
    const float *__restrict__ xrow = a.x[sc] + a.x_offset;

    // chunk c -> buffer b: the 160 samples before the chunk are its halo (for the span's first chunk they come
    // from the carry buffer instead and only the 8000 new samples are fetched)
    auto prefetch = [&](int c, int b) {
        const float *src = xrow + (size_t)c * kChunk16 - (c ? kHop : 0);
        float *dst = sm.xs[b] + (c ? 0 : kHop);
        const uint32_t bytes = (uint32_t)((c ? kXs : kChunk16) * sizeof(float));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic accesses to this buffer are ordered first
        tc::mbar_expect_tx(&sm.full[b], bytes);
        tc::bulk_g2s(dst, src, bytes, &sm.full[b]);
    };
    if (tid == 0) {
        tc::mbar_init(&sm.full[0], 1);
        tc::mbar_init(&sm.full[1], 1);
        tc::fence_mbar_init();
        prefetch(c_begin, 0);
    }
    for (int i = tid; i < kNfft; i += kStftThreads) {
        if (i < 256) sm.T[i] = gT[i];
        sm.win[i] = gwin[i];
    }
    __syncthreads();
    StftLaneConst lc;
    lc.init(sm.T, sm.win, lane);

    for (int c = c_begin, it = 0; c < c_end; ++c, ++it) {
        const int b = it & 1;
        tc::mbar_wait(&sm.full[b], (it >> 1) & 1);
        float *xs = sm.xs[b];
        if (c == 0 && tid < kHop) xs[tid] = a.halo_in[(size_t)sc * kHop + tid];
        // sum of squares of the chunk's own 8000 samples (same per-thread order as a strided 128-bit global read)
        float ss = 0.0f;
        for (int i = tid; i < kChunk16 / 4; i += kStftThreads) {
            const float4 v = *reinterpret_cast<const float4 *>(&xs[kHop + 4 * i]);
            ss = fmaf(v.x, v.x, ss);
            ss = fmaf(v.y, v.y, ss);
            ss = fmaf(v.z, v.z, ss);
            ss = fmaf(v.w, v.w, ss);
        }
        // The reduction's __syncthreads is the only block barrier of a chunk: every warp that passes it has finished the
        // frames of chunk c-1, so the other buffer is free and the prefetch of chunk c+1 goes out right behind it (a
        // whole chunk of transforms ahead of its use); it also makes the halo rows visible.  The partial sums alternate
        // between two rows so the next chunk's writes are a barrier away from this chunk's reads.
        stft_sumsq(a, ss, sm.red[b], sc, c, tid);
        if (tid == 0 && c + 1 < c_end) prefetch(c + 1, b ^ 1);
        if (c == a.K - 1 && tid < kHop) a.halo_out[(size_t)sc * kHop + tid] = xs[kFrames * kHop + tid];
        stft_frames(a, lc, xs, sm.zb[warp], sm.fh[warp], sc, c, warp, lane);
    }
}

// ---- D = 1, A/B option FVAD_STFT_STREAM=1: independent streaming warps (the ISTFT's structure) ----
// A warp owns one chunk: it reads each frame's 320 samples straight from global memory (the half shared with the
// previous frame hits L1), with the next frame's loads in flight while this one is transformed; the chunk's sum of
// squares accumulates over each frame's new hop.  No tile, no block barrier.
constexpr int kSsWarps = 8;
constexpr int kSsThreads = 32 * kSsWarps;

struct StftSmemS {
    float2 T[256];
    float win[kNfft];
    float2 zb[kSsWarps][kHop];
    __align__(16) __half fh[kSsWarps][176];
};

__global__ void __launch_bounds__(kSsThreads, 2) stft_stream_kernel(StftArgs a, const float2 *__restrict__ gT,
                                                                    const float *__restrict__ gwin) {
    __shared__ StftSmemS sm;
    const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    for (int i = tid; i < kNfft; i += kSsThreads) {
        if (i < 256) sm.T[i] = gT[i];
        sm.win[i] = gwin[i];
    }
    __syncthreads();
    const int item = blockIdx.x * kSsWarps + warp;
    if (item >= a.K * a.SC) return;
    const int sc = item / a.K, c = item - sc * a.K;
    StftLaneConst lc;
    lc.init(sm.T, sm.win, lane);
    float2 *Z = sm.zb[warp];
    __half *fh = sm.fh[warp];
    if (a.feat16 && lane < 15) fh[kBins + lane] = __float2half(0.0f);

    // frame f of the chunk = samples [160 (f-1), 160 (f-1) + 320) relative to the chunk start
    const float *__restrict__ xc = a.x[sc] + a.x_offset + (size_t)c * kChunk16;
    const float *__restrict__ halo = a.halo_in + (size_t)sc * kHop;
    float2 nx[5];
    auto fetch = [&](int f) {
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            const int j = lane + 32 * q;  // float2 index inside the frame; j < 80 is the hop shared with frame f-1
            const float *src = (c == 0 && f == 0 && j < kHop / 2) ? halo + 2 * j : xc + (f - 1) * kHop + 2 * j;
            nx[q] = __ldg(reinterpret_cast<const float2 *>(src));
        }
    };
    fetch(0);
    float ss = 0.0f;
    for (int f = 0; f < kFrames; ++f) {
        float2 v[5], raw[5];
#pragma unroll
        for (int q = 0; q < 5; ++q) raw[q] = nx[q];
        if (f + 1 < kFrames) fetch(f + 1);
#pragma unroll
        for (int q = 0; q < 5; ++q) {
            if (lane + 32 * q >= kHop / 2) {  // the frame's new hop: this chunk's own samples
                ss = fmaf(raw[q].x, raw[q].x, ss);
                ss = fmaf(raw[q].y, raw[q].y, ss);
            }
            v[q] = cscale(raw[q], lc.wreg[q]);
        }
        if (f == kFrames - 1 && c == a.K - 1) {  // the span's last hop is the next span's halo
#pragma unroll
            for (int q = 0; q < 5; ++q) {
                const int j = lane + 32 * q;
                if (j >= kHop / 2) reinterpret_cast<float2 *>(a.halo_out + (size_t)sc * kHop)[j - kHop / 2] = raw[q];
            }
        }
        stft_emit(a, lc, v, Z, fh, sc, c, f, lane);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (lane == 0) a.sumsq[(size_t)sc * (a.K + 1) + 1 + c] = ss;
}

// ---- general D (48 kHz input: D = 3): one chunk per CTA, decimating loads ----
struct StftSmemD {
    float xs[kXs];
    float2 T[256];
    float win[kNfft];
    float2 zb[kStftWarps][kHop];
    float red[kStftWarps];
    __align__(16) __half fh[kStftWarps][176];
};

template <int D>
__global__ void __launch_bounds__(kStftThreads) stft_kernel(StftArgs a, const float2 *__restrict__ gT,
                                                            const float *__restrict__ gwin) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    StftSmemD &sm = *reinterpret_cast<StftSmemD *>(smem_raw);
    float *xs = sm.xs;
    float2 *T = sm.T;  // indices used by the N=160 transform stay below 249
    float *win = sm.win;
    auto &zb = sm.zb;
    float *red = sm.red;
    auto &fh = sm.fh;

    const int c = blockIdx.x;    // chunk within the span
    const int sc = blockIdx.y;   // stream-channel
    const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // (tells the compiler the warp index is uniform)

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
        constexpr int kBatch = 5;  // loads in flight per thread
        for (int i0 = tid; i0 < n4; i0 += kStftThreads * kBatch) {
            float4 v[kBatch];
#pragma unroll
            for (int u = 0; u < kBatch; ++u) {
                const int i = i0 + u * kStftThreads;
                v[u] = i < n4 ? __ldg(x4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < kBatch; ++u) {
                const int i = i0 + u * kStftThreads;
                if (i >= n4) break;
                ss = fmaf(v[u].x, v[u].x, ss);
                ss = fmaf(v[u].y, v[u].y, ss);
                ss = fmaf(v[u].z, v[u].z, ss);
                ss = fmaf(v[u].w, v[u].w, ss);
                // sample index s = 4i+j is kept when s % D == 0, at xs[hop + s/D]
                const int s0 = 4 * i;
                const float e[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if ((s0 + j) % D == 0) xs[kHop + (s0 + j) / D] = e[j];
            }
        }
    }
    // halo: the 160 decimated samples that precede this chunk
    if (tid < kHop) xs[tid] = (c == 0) ? a.halo_in[(size_t)sc * kHop + tid] : __ldg(x - (size_t)(kHop - tid) * D);
    stft_sumsq(a, ss, red, sc, c, tid);
    // the last chunk of the span leaves its last hop as the next span's halo
    if (c == a.K - 1 && tid < kHop) a.halo_out[(size_t)sc * kHop + tid] = xs[kFrames * kHop + tid];

    StftLaneConst lc;
    lc.init(T, win, lane);
    stft_frames(a, lc, xs, zb[warp], fh[warp], sc, c, warp, lane);
}

void launch_stft(const StftArgs &a, const Tables &t, cudaStream_t s) {
    if (a.K <= 0 || a.SC <= 0) return;
    static const bool stream_variant = getenv("FVAD_STFT_STREAM") != nullptr;
    if (a.D == 1 && stream_variant) {
        const int items = a.K * a.SC;
        stft_stream_kernel<<<(unsigned)((items + kSsWarps - 1) / kSsWarps), kSsThreads, 0, s>>>(a, t.T320, t.win320);
    } else if (a.D == 1) {
        static bool attr = false;
        if (!attr) {
            FVAD_CUDA_CHECK(cudaFuncSetAttribute(stft_kernel_d1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(StftSmem1)));
            attr = true;
        }
        dim3 grid((unsigned)((a.K + kChunksPerCta - 1) / kChunksPerCta), (unsigned)a.SC);
        stft_kernel_d1<<<grid, kStftThreads, sizeof(StftSmem1), s>>>(a, t.T320, t.win320);
    } else if (a.D == 3) {
        static bool attr3 = false;
        if (!attr3) {
            FVAD_CUDA_CHECK(cudaFuncSetAttribute(stft_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(StftSmemD)));
            attr3 = true;
        }
        dim3 grid((unsigned)a.K, (unsigned)a.SC);
        stft_kernel<3><<<grid, kStftThreads, sizeof(StftSmemD), s>>>(a, t.T320, t.win320);
    } else {
        throw Error(FVAD_ERR_INVALID_SAMPLE_RATE, "unsupported decimation factor");
    }
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace fvad
