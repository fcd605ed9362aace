// nn_tc_gru.cu -- one GRU layer (400 units, 54 steps, zero initial state) as a persistent tcgen05
// kernel: a CTA owns 128 sequences for all 54 steps; the hidden state never leaves the SM.
//
// Replaces the ONNX GRU nodes inside `onnx_instance.run()` (reference src/NSNet2.zig:220; ONNX GRU
// semantics with linear_before_reset = 1, gate order z,r,h -- SURVEY App. C):
//     z = s(xz + Rz h + b), r = s(xr + Rr h + b), n = tanh(xn + r*(Rn h + Rbn)), h' = (1-z)*n + z*h
//
// Data flow per step: h_{t-1} (fp16) sits in TENSOR MEMORY as the A operand (128 lanes x 200
// columns, two halves per column); the recurrent weights R stream from L2 through a shared-memory
// ring in 25 chunks of 48 gate rows (16 units x {z,r,n}); tcgen05.mma (A from TMEM) writes each
// chunk's 128x48 fp32 pre-activations into one of two TMEM accumulators; two epilogue warpgroups
// (even / odd chunks) read them back, add the precomputed input projection xp, apply the gates and
// write h_t both into the other TMEM h buffer (next step's A operand) and to global memory (next
// layer's input).  TMEM columns: h ping [0,200) | h pong [200,400) | acc0 [400,448) | acc1 [448,496).
#include "nn_tc.cuh"
#include "tc_dev.cuh"

namespace fvad {
namespace tc {

namespace {

// timeline probe: event e of (step t, chunk c) of the first tile of CTA 0
#define GRU_TRACE(e, t, c) do { if (a.trace && blockIdx.x == 0 && iq == 0) a.trace[((t) * 25 + (c)) * 8 + (e)] = (unsigned long long)clock64(); } while (0)

constexpr int kH = 400;
constexpr int kChunkUnits = 16;
constexpr int kChunkRows = 48;
constexpr int kChunks = kH / kChunkUnits;              // 25
constexpr int kRChunkBytes = kChunkRows * kH * 2;      // 38 400
constexpr int kRStages = 5;
constexpr int kGruThreads = 384;                       // 4 control warps + 2 x 4 epilogue warps
constexpr int kGruSmem = kRStages * kRChunkBytes + kH * 4 + 1024;
constexpr uint32_t kColH0 = 0, kColH1 = 200, kColAcc0 = 400, kColAcc1 = 448;

struct GruShared {
    uint64_t full[kRStages], empty[kRStages], acc_full[2], acc_empty[2], h_ready;
    uint32_t tmem_base;
};

// Gate activations on the hardware tanh (MUFU.TANH, tanh.approx.f32: relative error ~2^-11, the rounding
// step of the fp16 hidden state these gates feed): sigma(x) = 0.5 + 0.5*tanh(x/2).  One special-function op per
// gate instead of ex2 + rcp -- the epilogue warps, not the tensor pipe, bound this kernel (ncu: XU pipe and
// issue slots), so this halves their work.  The fp32 backend (nn_f32.cu) keeps expf/tanhf.
__device__ __forceinline__ float tanh_hw(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// (the argument arrives HALVED: nn_tc.cu stores the z and r rows of W, R and their biases times 1/2)
__device__ __forceinline__ float sigmoid_fast(float half_x) { return fmaf(0.5f, tanh_hw(half_x), 0.5f); }
__device__ __forceinline__ float tanh_fast(float x) { return tanh_hw(x); }

__global__ void __launch_bounds__(kGruThreads, 1) tc_gru_kernel(const GruArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    float *rbn_s = reinterpret_cast<float *>(smem + kRStages * kRChunkBytes);
    GruShared *sh = reinterpret_cast<GruShared *>(smem + kRStages * kRChunkBytes + kH * 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kRStages; ++i) {
            mbar_init(&sh->full[i], 1);
            mbar_init(&sh->empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sh->acc_full[i], 1);
            mbar_init(&sh->acc_empty[i], 4);
        }
        mbar_init(&sh->h_ready, 8);
        fence_mbar_init();
    }
    for (int i = threadIdx.x; i < kH; i += kGruThreads) rbn_s[i] = a.rbn[i];
    if (warp == 2) tmem_alloc<512>(&sh->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sh->tmem_base;

    const int n_q = (a.Q - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // tiles this CTA processes

    if (warp == 0) {
        // ===== producer: R chunks, identical sequence every step =====
        if (elect_one()) {
            uint32_t it = 0;
            const uint8_t *R = reinterpret_cast<const uint8_t *>(a.Rimg);
            for (int iq = 0; iq < n_q; ++iq)
                for (int t = 0; t < kSteps; ++t)
                    for (int c = 0; c < kChunks; ++c, ++it) {
                        const int st = it % kRStages;
                        mbar_wait(&sh->empty[st], ((it / kRStages) & 1) ^ 1);
                        mbar_expect_tx(&sh->full[st], kRChunkBytes);
                        bulk_g2s(smem + st * kRChunkBytes, R + (size_t)c * kRChunkBytes, kRChunkBytes, &sh->full[st]);
                    }
        }
    } else if (warp == 1 || warp == 3) {
        // ===== MMA issuers: warp 1 issues the even chunks (accumulator 0), warp 3 the odd ones (accumulator 1) =====
        // One issuing thread spends ~520 cycles pushing a chunk's 25 MMAs and ~400 more on its two barrier waits,
        // commits and loop overhead (timeline: tools/gru_trace.py) -- alone it leaves the tensor pipe idle a third of
        // the time.  Two threads alternate, so one's bookkeeping hides under the other's issue burst.  The chunks
        // are independent (separate accumulators, separate R stages); each thread's commits cover its own MMAs.
        if (elect_one()) {
            constexpr uint32_t idesc = make_idesc_f16(kTileM, kChunkRows);
            const uint32_t g = warp == 1 ? 0u : 1u;
            const uint32_t d = tmem + (g ? kColAcc1 : kColAcc0);
            uint32_t step = 0, use = 0;
            for (int iq = 0; iq < n_q; ++iq)
                for (int t = 0; t < kSteps; ++t, ++step) {
                    // h_{t-1} complete in TMEM (epilogue warps of the previous step / the zero fill)
                    mbar_wait(&sh->h_ready, step & 1);
                    tc_fence_after();
                    if (g == 0) GRU_TRACE(6, t, 0);
                    const uint32_t hold = tmem + ((t & 1) ? kColH1 : kColH0);
                    for (int c = (int)g; c < kChunks; c += 2) {
                        const uint32_t it = step * kChunks + (uint32_t)c;
                        const uint32_t st = it % kRStages;
                        mbar_wait(&sh->full[st], (it / kRStages) & 1);
                        mbar_wait(&sh->acc_empty[g], (use & 1) ^ 1);
                        ++use;
                        tc_fence_after();
                        GRU_TRACE(1, t, c);
                        // K-step kk reads h columns [8kk, 8kk+8) and the R slab 1536 bytes (96 descriptor units) further
                        const uint64_t bd0 = make_smem_desc(smem_u32(smem + st * kRChunkBytes), kChunkRows * 16, 128);
                        const uint32_t blo = (uint32_t)bd0, bhi = (uint32_t)(bd0 >> 32);
                        mma_ts_lohi<false>(d, hold, blo, bhi, idesc);
#pragma unroll
                        for (int kk = 1; kk < kH / 16; ++kk)
                            mma_ts_lohi<true>(d, hold + kk * 8, blo + kk * (2 * kChunkRows), bhi, idesc);
                        mma_commit(&sh->empty[st]);
                        mma_commit(&sh->acc_full[g]);
                        GRU_TRACE(2, t, c);
                    }
                }
        }
    } else if (warp >= 4) {
        // ===== epilogue: group g = even / odd chunks =====
        const int g = (warp - 4) >> 2;
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        const uint32_t lane_off = (uint32_t)(wq * 32) << 16;
        const uint32_t acc = tmem + lane_off + (g ? kColAcc1 : kColAcc0);
        const size_t row_off = (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16;
        uint32_t use = 0;
        for (int iq = 0; iq < n_q; ++iq) {
            const int q = blockIdx.x + iq * gridDim.x;
            // h_{-1} = 0: zero this group's columns of the ping buffer
            {
                uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                for (int c = g; c < kChunks; c += 2) tmem_st8(tmem + lane_off + kColH0 + c * 8, z);
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&sh->h_ready);
            }
            for (int t = 0; t < kSteps; ++t) {
                const uint32_t hold = tmem + lane_off + ((t & 1) ? kColH1 : kColH0);
                const uint32_t hnew = tmem + lane_off + ((t & 1) ? kColH0 : kColH1);
                const uint8_t *xpt = reinterpret_cast<const uint8_t *>(a.xp) + (size_t)(q * kSteps + t) * tile_bytes(3 * kH) + row_off;
                uint8_t *hot = reinterpret_cast<uint8_t *>(a.hout) + (size_t)(q * kSteps + t) * tile_bytes(kH) + row_off;
                const bool store = t >= a.t_store_min;
                uint4 xn[6];
#pragma unroll
                for (int i = 0; i < 6; ++i) xn[i] = __ldg(reinterpret_cast<const uint4 *>(xpt + (size_t)(g * 6 + i) * (kTileM * 16)));
                for (int c = g; c < kChunks; c += 2) {
                    uint4 x[6];
#pragma unroll
                    for (int i = 0; i < 6; ++i) x[i] = xn[i];
                    if (c + 2 < kChunks) {
#pragma unroll
                        for (int i = 0; i < 6; ++i)
                            xn[i] = __ldg(reinterpret_cast<const uint4 *>(xpt + (size_t)((c + 2) * 6 + i) * (kTileM * 16)));
                    }
                    mbar_wait(&sh->acc_full[g], use & 1);
                    ++use;
                    tc_fence_after();
                    if (wq == 0 && lane == 0) GRU_TRACE(3, t, c);
                    uint32_t az[16], ar[16], an[16], ho[8];
                    tmem_ld16(acc, az);
                    tmem_ld16(acc + 16, ar);
                    tmem_ld16(acc + 32, an);
                    tmem_ld8(hold + c * 8, ho);
                    tmem_wait_ld();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sh->acc_empty[g]);  // accumulator drained: the next chunk's MMAs may start
                    if (wq == 0 && lane == 0) GRU_TRACE(4, t, c);

                    const uint32_t *xw = reinterpret_cast<const uint32_t *>(x);  // 24 words: z[0..7] r[8..15] n[16..23] (pairs)
                    uint32_t hn[8];
#pragma unroll
                    for (int p = 0; p < 8; ++p) {
                        const float2 xz = unpack_h2(xw[p]), xr = unpack_h2(xw[8 + p]), xnn = unpack_h2(xw[16 + p]);
                        const float2 hp = unpack_h2(ho[p]);
                        const float b0 = rbn_s[c * kChunkUnits + 2 * p], b1 = rbn_s[c * kChunkUnits + 2 * p + 1];
                        const float z0 = sigmoid_fast(xz.x + __uint_as_float(az[2 * p]));
                        const float z1 = sigmoid_fast(xz.y + __uint_as_float(az[2 * p + 1]));
                        const float r0 = sigmoid_fast(xr.x + __uint_as_float(ar[2 * p]));
                        const float r1 = sigmoid_fast(xr.y + __uint_as_float(ar[2 * p + 1]));
                        const float n0 = tanh_fast(xnn.x + r0 * (__uint_as_float(an[2 * p]) + b0));
                        const float n1 = tanh_fast(xnn.y + r1 * (__uint_as_float(an[2 * p + 1]) + b1));
                        const float h0 = n0 + z0 * (hp.x - n0);
                        const float h1 = n1 + z1 * (hp.y - n1);
                        hn[p] = pack_h2(h0, h1);
                    }
                    tmem_st8(hnew + c * 8, hn);
                    if (wq == 0 && lane == 0) GRU_TRACE(5, t, c);
                    if (store) {
                        uint8_t *p = hot + (size_t)(c * 2) * (kTileM * 16);
                        *reinterpret_cast<uint4 *>(p) = make_uint4(hn[0], hn[1], hn[2], hn[3]);
                        *reinterpret_cast<uint4 *>(p + kTileM * 16) = make_uint4(hn[4], hn[5], hn[6], hn[7]);
                    }
                }
                // h_t complete for this warp's rows and this group's columns
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0 && t < kSteps - 1) mbar_arrive(&sh->h_ready);  // the next tile's zero fill arrives for its t = 0
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}

}  // namespace

void launch_tc_gru(const GruArgs &a, int n_sm, cudaStream_t s) {
    static bool attr = false;
    if (!attr) {
        FVAD_CUDA_CHECK(cudaFuncSetAttribute(tc_gru_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGruSmem));
        attr = true;
    }
    if (a.Q <= 0) return;
    const int grid = a.Q < n_sm ? a.Q : n_sm;
    tc_gru_kernel<<<grid, kGruThreads, kGruSmem, s>>>(a);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace tc
}  // namespace fvad
