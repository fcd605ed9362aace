// istft.cu -- K3a: gain clamp x spectrum -> inverse real FFT-320 -> window/320 -> overlap-add.
//
// Replaces, per chunk of one channel (reference src/NSNet2.zig:221-229):
//   NSNet2.applySpecgramGain   (NSNet2.zig:289-310)  g = clamp(gain, -80, 1); re *= g; im *= g
//   NSNet2.reconstructAudio    (NSNet2.zig:312-339) -> FFT.invFft (FFT.zig:115-134) -> kiss_fftri,
//                              then *= window[i] * (1/320) and += into audio_output
// and the output-halo carry of NSNet2.denoise (NSNet2.zig:181-184,197,201).
//
// Default kernel (istft_stream_kernel): every WARP owns a run of L consecutive frames of one
// stream-channel and is independent of every other warp -- no block barrier, no shared output tile.
// After the inverse transform lane l, register k1 holds output pair n = k1 + 5*bitrev5(l): odd lanes
// hold the frame's second half (its tail), even lanes the first half (its head), and lane l^1 holds
// pair n +- 80.  So one shfl_xor(1) hands the tail of frame g to the lanes that will hold the head of
// frame g+1, the overlap-add happens in registers (head + tail: the same two-term IEEE sum the
// reference forms with +=), and a finished 160-sample hop leaves through a 640-byte per-warp staging
// row as coalesced 16-byte stores.  The spectrum and gain rows of frame g+1 are loaded into registers
// before frame g is transformed, so the global-load latency hides behind the FFT.  A run that does
// not start the span first inverts the frame before it for its tail (1/L redundant work; L is
// chosen on the host so the runs fill whole waves of the grid).
//
// A/B option FVAD_ISTFT_CHUNK=1 (istft_chunk_kernel, the earlier version): one CTA per (chunk,
// stream-channel) inverts 51 frames into a shared tile; even local frames store, then odd ones add.
// ncu showed half of its stall samples on the row loads and 40 % on the two block barriers.
// Output is the 16 kHz denoised stream y16.
#include <cstdlib>

#include "fft_dev.cuh"
#include "kernels.cuh"

namespace fvad {

constexpr int kIstftThreads = 320;
constexpr int kIstftWarps = kIstftThreads / 32;
constexpr int kLocalFrames = kFrames + 1;  // local frame lf <-> chunk frame lf-1

__global__ void __launch_bounds__(kIstftThreads) istft_chunk_kernel(IstftArgs a, const float2 *__restrict__ gT,
                                                              const float *__restrict__ gwn) {
    __shared__ __align__(16) float y[(kLocalFrames + 1) * kHop];  // 52 hops
    __shared__ float2 T[256];
    __shared__ float2 xb[kIstftWarps][kBins + 1];

    const int c = blockIdx.x, sc = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // (tells the compiler the warp index is uniform)
    for (int i = tid; i < 256; i += kIstftThreads) T[i] = gT[i];
    const float *__restrict__ wn = gwn;  // 1.25 KB, L1-resident; keeps static shared memory under 48 KB
    __syncthreads();

    // lane-constant operands in registers for all frames of this warp: FFT twiddles, the pre-pack
    // twiddles i*conj(T[k]) and the synthesis window (window*1/320) at this lane's output positions
    WarpFftTw<5, true> ftw;
    ftw.init(T, lane);
    float2 itw[5], wreg[5];
#pragma unroll
    for (int q = 0; q < 5; ++q) {
        const float2 t = T[lane + 32 * q];
        itw[q] = make_float2(t.y, t.x);
        wreg[q] = __ldg(reinterpret_cast<const float2 *>(wn) + cfft_out_index<5>(lane, q));
    }
    float2 *X = xb[warp];
    const size_t row0 = (size_t)sc * (kFrames * a.K) + (size_t)c * kFrames;  // row of chunk frame 0
    for (int pass = 0; pass < 2; ++pass) {
        for (int lf = 2 * warp + pass; lf < kLocalFrames; lf += 2 * kIstftWarps) {
            float *yo = y + lf * kHop;
            if (lf == 0 && c == 0) {
                // frame before the span: only its tail matters, and it comes from the carry
                for (int i = lane; i < kHop; i += 32) {
                    yo[i] = 0.0f;
                    yo[kHop + i] = a.tail_in[(size_t)sc * kHop + i];
                }
                continue;
            }
            const size_t row = row0 + lf - 1;
            const float2 *sp = reinterpret_cast<const float2 *>(a.spec) + row * kBins;
            const float *gp = a.gains + row * kBins;
#pragma unroll
            for (int q = 0; q < 6; ++q) {
                const int k = lane + 32 * q;
                if (k <= kHop) {
                    float g = __ldg(gp + k);
                    g = g < -80.0f ? -80.0f : (g > 1.0f ? 1.0f : g);  // NSNet2.zig:295-305
                    const float2 s = __ldg(sp + k);
                    X[k] = make_float2(s.x * g, s.y * g);
                }
            }
            __syncwarp();
            float2 v[5];
#pragma unroll
            for (int q = 0; q < 5; ++q) {  // kiss_fftri pre-pack (kiss_fftr.c:131-153)
                const int k = lane + 32 * q;
                const float2 xa = X[k], xb2 = X[kHop - k];
                const float2 e = make_float2(xa.x + xb2.x, xa.y - xb2.y), d = make_float2(xa.x - xb2.x, xa.y + xb2.y);
                v[q] = cadd(e, cmul(d, itw[q]));
            }
            warp_cfft_tw<5, true>(v, ftw);
#pragma unroll
            for (int k1 = 0; k1 < 5; ++k1) {
                const int n = cfft_out_index<5>(lane, k1);
                // NSNet2.zig:335-336: inv *= window*norm; out += inv
                float2 r = make_float2(v[k1].x * wreg[k1].x, v[k1].y * wreg[k1].y);
                float2 *dst = reinterpret_cast<float2 *>(yo) + n;
                if (pass == 1) {
                    const float2 o = *dst;
                    r.x += o.x;
                    r.y += o.y;
                }
                *dst = r;
            }
            __syncwarp();
        }
        __syncthreads();
    }
    // chunk output = local hops 1..50
    float *out = a.y16 + (size_t)sc * (kYLook + (size_t)kChunk16 * a.K) + kYLook + (size_t)c * kChunk16;
    for (int i = tid; i < kChunk16 / 4; i += kIstftThreads)
        reinterpret_cast<float4 *>(out)[i] = *reinterpret_cast<const float4 *>(&y[kHop + 4 * i]);
    // tail of the span's last frame -> carry for the next span
    if (c == a.K - 1 && tid < kHop) a.tail_out[(size_t)sc * kHop + tid] = y[kLocalFrames * kHop + tid];
}

// ---- streaming warps ----
constexpr int kIsWarps = 8;
constexpr int kIsThreads = 32 * kIsWarps;
// CTAs per SM (A/B with FVAD_ISTFT_CTAS, ms per 20 stream-hours on one box): 2 -> 117 registers, 5.17; 3 -> 80 registers,
// 5.89; 4 -> 64 registers with 132 B of spills, 7.5.  With room for the 18 prefetch registers the next frame's loads
// issue a whole transform ahead; more resident warps do not make up for losing that.
constexpr int kIsCtasPerSm = 2;

template <int CTAS>
__global__ void __launch_bounds__(kIsThreads, CTAS) istft_stream_kernel(IstftArgs a, const float2 *__restrict__ gT,
                                                                                const float *__restrict__ gwn, int L,
                                                                                int runs_per_sc, int total_runs) {
    __shared__ float2 T[256];
    __shared__ float2 xb[kIsWarps][kBins + 1];
    __shared__ __align__(16) float hopb[kIsWarps][kHop];

    const int tid = threadIdx.x, lane = tid & 31, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    for (int i = tid; i < 256; i += kIsThreads) T[i] = gT[i];
    __syncthreads();
    const int run = blockIdx.x * kIsWarps + warp;
    if (run >= total_runs) return;
    const int sc = run / runs_per_sc, r = run - sc * runs_per_sc;
    const int rows = kFrames * a.K;
    const int g0 = r * L, g1 = min(rows, g0 + L);

    WarpFftTw<5, true> ftw;
    ftw.init(T, lane);
    float2 itw[5], wreg[5];
#pragma unroll
    for (int q = 0; q < 5; ++q) {
        const float2 t = T[lane + 32 * q];
        itw[q] = make_float2(t.y, t.x);
        wreg[q] = __ldg(reinterpret_cast<const float2 *>(gwn) + cfft_out_index<5>(lane, q));
    }
    const bool even = (lane & 1) == 0;
    const int nb = 5 * bitrev5(lane);  // even lanes: nb < 80
    float2 *X = xb[warp];
    float *hb = hopb[warp];
    const float2 *sp = reinterpret_cast<const float2 *>(a.spec) + (size_t)sc * rows * kBins;
    const float *gp = a.gains + (size_t)sc * rows * kBins;
    float *out = a.y16 + (size_t)sc * (kYLook + (size_t)kChunk16 * a.K) + kYLook;

    // tail of the frame before the run, already in the lanes that will add it
    float2 carry[5];
    int g = g0 - 1;
    if (r == 0) {
        g = g0;
#pragma unroll
        for (int k1 = 0; k1 < 5; ++k1)
            carry[k1] = even ? __ldg(reinterpret_cast<const float2 *>(a.tail_in + (size_t)sc * kHop) + nb + k1)
                             : make_float2(0.0f, 0.0f);
    }
    float2 ps[6];
    float pg[6];
    auto fetch = [&](int f) {
        const float2 *s = sp + (size_t)f * kBins;
        const float *gg = gp + (size_t)f * kBins;
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            const int k = lane + 32 * q;
            if (q < 5 || k <= kHop) {
                ps[q] = __ldg(s + k);
                pg[q] = __ldg(gg + k);
            }
        }
    };
    fetch(g);
    for (; g < g1; ++g) {
#pragma unroll
        for (int q = 0; q < 6; ++q) {
            const int k = lane + 32 * q;
            if (q < 5 || k <= kHop) {
                float gn = pg[q];
                gn = gn < -80.0f ? -80.0f : (gn > 1.0f ? 1.0f : gn);  // NSNet2.zig:295-305
                X[k] = make_float2(ps[q].x * gn, ps[q].y * gn);
            }
        }
        if (g + 1 < g1) fetch(g + 1);
        __syncwarp();
        float2 v[5];
#pragma unroll
        for (int q = 0; q < 5; ++q) {  // kiss_fftri pre-pack (kiss_fftr.c:131-153)
            const int k = lane + 32 * q;
            const float2 xa = X[k], xb2 = X[kHop - k];
            const float2 e = cfma(xb2, make_float2(1.0f, -1.0f), xa), d = cfma(xb2, make_float2(-1.0f, 1.0f), xa);  // exact products
            v[q] = cadd(e, cmul(d, itw[q]));
        }
        warp_cfft_tw<5, true>(v, ftw);
#pragma unroll
        for (int k1 = 0; k1 < 5; ++k1)  // NSNet2.zig:335-336: inv *= window*norm
            v[k1] = cscale(v[k1], wreg[k1]);  // (a packed multiply rounds each component: nothing to contract with)
        if (g >= g0 && even) {
#pragma unroll
            for (int k1 = 0; k1 < 5; ++k1)  // out += inv: this frame's head + the previous frame's tail
                reinterpret_cast<float2 *>(hb)[nb + k1] = cadd(v[k1], carry[k1]);
        }
#pragma unroll
        for (int k1 = 0; k1 < 5; ++k1) {
            carry[k1].x = __shfl_xor_sync(0xffffffffu, v[k1].x, 1);
            carry[k1].y = __shfl_xor_sync(0xffffffffu, v[k1].y, 1);
        }
        __syncwarp();
        if (g >= g0) {
            float4 *o4 = reinterpret_cast<float4 *>(out + (size_t)g * kHop);
            const float4 *h4 = reinterpret_cast<const float4 *>(hb);
            o4[lane] = h4[lane];
            if (lane < kHop / 4 - 32) o4[32 + lane] = h4[32 + lane];
        }
        __syncwarp();
    }
    // tail of the span's last frame -> carry for the next span
    if (g1 == rows && even) {
#pragma unroll
        for (int k1 = 0; k1 < 5; ++k1)
            reinterpret_cast<float2 *>(a.tail_out + (size_t)sc * kHop)[nb + k1] = carry[k1];
    }
}

// run length: the one in [24, 64] that wastes least on the partial last wave and on the 1/L redundant frame
static int pick_run_length(int rows, int SC, int ctas) {
    int sms = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const double slots = (double)sms * ctas * kIsWarps;
    int best = 40;
    double best_cost = 1e30;
    for (int L = 24; L <= 64; ++L) {
        const double runs = (double)SC * ((rows + L - 1) / L);
        const double waves = (double)(long long)((runs + slots - 1) / slots);
        const double cost = waves * (L + 1);  // frames inverted by the busiest warp slot
        if (cost < best_cost) best_cost = cost, best = L;
    }
    return best;
}

void launch_istft(const IstftArgs &a, const Tables &t, cudaStream_t s) {
    if (a.K <= 0 || a.SC <= 0) return;
    static const bool chunk_variant = getenv("FVAD_ISTFT_CHUNK") != nullptr;
    if (!chunk_variant) {
        const int rows = kFrames * a.K;
        thread_local int cached_rows = -1, cached_sc = -1, cached_L = 0;  // (any L is correct; the cache only saves the search)
        static const int ctas = getenv("FVAD_ISTFT_CTAS") ? atoi(getenv("FVAD_ISTFT_CTAS")) : kIsCtasPerSm;
        if (rows != cached_rows || a.SC != cached_sc) cached_L = pick_run_length(rows, a.SC, ctas), cached_rows = rows, cached_sc = a.SC;
        const int L = cached_L;
        const int runs_per_sc = (rows + L - 1) / L, total = runs_per_sc * a.SC;
        const unsigned grid = (unsigned)((total + kIsWarps - 1) / kIsWarps);
        if (ctas == 4) istft_stream_kernel<4><<<grid, kIsThreads, 0, s>>>(a, t.T320, t.win320n, L, runs_per_sc, total);
        else if (ctas == 3) istft_stream_kernel<3><<<grid, kIsThreads, 0, s>>>(a, t.T320, t.win320n, L, runs_per_sc, total);
        else istft_stream_kernel<kIsCtasPerSm><<<grid, kIsThreads, 0, s>>>(a, t.T320, t.win320n, L, runs_per_sc, total);
        FVAD_LAUNCHED();
        FVAD_CUDA_CHECK(cudaGetLastError());
        return;
    }
    dim3 grid((unsigned)a.K, (unsigned)a.SC);
    istft_chunk_kernel<<<grid, kIstftThreads, 0, s>>>(a, t.T320, t.win320n);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace fvad
