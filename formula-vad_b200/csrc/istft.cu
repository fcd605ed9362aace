// istft.cu -- K3a: gain clamp x spectrum -> inverse real FFT-320 -> window/320 -> overlap-add.
//
// Replaces, per chunk of one channel (reference src/NSNet2.zig:221-229):
//   NSNet2.applySpecgramGain   (NSNet2.zig:289-310)  g = clamp(gain, -80, 1); re *= g; im *= g
//   NSNet2.reconstructAudio    (NSNet2.zig:312-339) -> FFT.invFft (FFT.zig:115-134) -> kiss_fftri,
//                              then *= window[i] * (1/320) and += into audio_output
// and the output-halo carry of NSNet2.denoise (NSNet2.zig:181-184,197,201).
//
// One CTA per (chunk, stream-channel).  The chunk's 8000 output samples receive the heads of its
// own 50 frames and the tails of frames -1..48, so the CTA inverts 51 frames (frame -1 is the
// previous chunk's last frame; at the span start its tail comes from the carry buffer).  Frames
// with even local index store, then odd ones add: no atomics, and the two-term sum is the same
// IEEE addition the reference performs.  Output is the 16 kHz denoised stream y16.
#include "fft_dev.cuh"
#include "kernels.cuh"

namespace fvad {

constexpr int kIstftThreads = 320;
constexpr int kIstftWarps = kIstftThreads / 32;
constexpr int kLocalFrames = kFrames + 1;  // local frame lf <-> chunk frame lf-1

__global__ void __launch_bounds__(kIstftThreads) istft_kernel(IstftArgs a, const float2 *__restrict__ gT,
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

void launch_istft(const IstftArgs &a, const Tables &t, cudaStream_t s) {
    if (a.K <= 0 || a.SC <= 0) return;
    dim3 grid((unsigned)a.K, (unsigned)a.SC);
    istft_kernel<<<grid, kIstftThreads, 0, s>>>(a, t.T320, t.win320n);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace fvad
