// nn_tc_gru2.cu -- the GRU layer of nn_tc_gru.cu on CTA PAIRS (tcgen05 cta_group::2).
//
// Why pairs: in the one-CTA kernel every byte of the recurrent weights R is written into shared memory by
// the bulk copy and read back out of it by the tensor core once per 128 sequences.  At the tensor-pipe
// rate that is 64 B/clk in + 64 B/clk out, the whole 128 B/clk of an SM's shared memory, and the kernel
// settles at ~50 % tensor-pipe activity with the MMA queue back-pressured (ncu: stall_mio on UTCHMMA,
// profiles/r01_*).  A pair of CTAs (two SMs of one cluster) runs ONE 256-row MMA per K-step: each CTA keeps
// its own 128 sequences' hidden state and accumulators in its own TMEM, and only HALF of each R chunk
// (24 of the 48 gate rows) in its own shared memory; the tensor cores of both SMs read both halves.
// Shared-memory traffic per SM halves, so does the L2 read stream.
//
// Same math and data layout as nn_tc_gru.cu (see there for the gate equations, reference
// src/NSNet2.zig:220 / ONNX GRU linear_before_reset = 1).  Differences:
//   * cluster (2,1,1); CTA rank r owns tile q = 2*pair + r; only rank 0 issues MMAs and commits (multicast
//     to both CTAs' barriers);
//   * R image: per chunk two half images of 24 rows (tiled with R = 24), rank r fetches half r;
//   * the leader must know the peer's half has landed: a relay thread in the peer forwards its `full`
//     barrier to the leader's `peer_full`;
//   * accumulator-drained and h-ready arrivals of the peer's epilogue warps go to the leader's barriers
//     (remote mbarrier.arrive, release.cluster).
#include "nn_tc.cuh"
#include "tc_dev.cuh"

namespace fvad {
namespace tc {

namespace {

constexpr int kH = 400;
constexpr int kChunkUnits = 16;
constexpr int kChunkRows = 48;                      // N of the pair's MMA
constexpr int kHalfRows = kChunkRows / 2;           // B rows held per CTA
constexpr int kChunks = kH / kChunkUnits;           // 25
constexpr int kHalfBytes = kHalfRows * kH * 2;      // 19 200
constexpr int kStages = 8;
constexpr int kGru2Threads = 384;                   // 4 control warps + 2 x 4 epilogue warps
constexpr int kGru2Smem = kStages * kHalfBytes + kH * 4 + 1024;
constexpr uint32_t kColH0 = 0, kColH1 = 200, kColAcc0 = 400, kColAcc1 = 448;

struct Gru2Shared {
    uint64_t full[kStages], peer_full[kStages], empty[kStages], acc_full[2], acc_empty[2], h_ready;
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
__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(0.5f, tanh_hw(0.5f * x), 0.5f); }
__device__ __forceinline__ float tanh_fast(float x) { return tanh_hw(x); }

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGru2Threads, 1) tc_gru2_kernel(const GruArgs a, int n_pairs_total) {
    extern __shared__ __align__(1024) uint8_t smem[];
    float *rbn_s = reinterpret_cast<float *>(smem + kStages * kHalfBytes);
    Gru2Shared *sh = reinterpret_cast<Gru2Shared *>(smem + kStages * kHalfBytes + kH * 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair0 = blockIdx.x >> 1, n_pair_ctas = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&sh->full[i], 1);
            mbar_init(&sh->peer_full[i], 1);
            mbar_init(&sh->empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sh->acc_full[i], 1);
            mbar_init(&sh->acc_empty[i], 8);   // 4 epilogue warps of the group in each CTA
        }
        mbar_init(&sh->h_ready, 16);           // 8 epilogue warps in each CTA
        fence_mbar_init();
    }
    for (int i = threadIdx.x; i < kH; i += kGru2Threads) rbn_s[i] = a.rbn[i];
    if (warp == 2) tmem_alloc2<512>(&sh->tmem_base);
    tc_fence_before();
    __syncthreads();
    cluster_sync();  // both CTAs' barriers and TMEM exist before any remote arrive / paired MMA
    tc_fence_after();
    const uint32_t tmem = sh->tmem_base;

    const int n_p = (n_pairs_total - pair0 + n_pair_ctas - 1) / n_pair_ctas;  // tile pairs this cluster processes
    const uint32_t n_fill = (uint32_t)n_p * kSteps * kChunks;

    if (warp == 0) {
        // ===== producer: this CTA's half of every R chunk, identical sequence every step =====
        if (elect_one()) {
            const uint8_t *R = reinterpret_cast<const uint8_t *>(a.Rimg) + (size_t)rank * kHalfBytes;
            uint32_t it = 0;
            for (int ip = 0; ip < n_p; ++ip)
                for (int t = 0; t < kSteps; ++t)
                    for (int c = 0; c < kChunks; ++c, ++it) {
                        const int st = it % kStages;
                        mbar_wait(&sh->empty[st], ((it / kStages) & 1) ^ 1);
                        mbar_expect_tx(&sh->full[st], kHalfBytes);
                        bulk_g2s(smem + st * kHalfBytes, R + (size_t)c * (2 * kHalfBytes), kHalfBytes, &sh->full[st]);
                    }
            // drain: the leader's multicast commits to this CTA's `empty` barriers must have landed before exit
            for (uint32_t s = 0; s < (uint32_t)kStages && s < n_fill; ++s) {
                const uint32_t last = ((n_fill - 1 - s) / kStages) * kStages + s;  // last fill of stage s
                mbar_wait(&sh->empty[s], (last / kStages) & 1);
            }
        }
    } else if (warp == 3) {
        // ===== relay (peer only): tell the leader that this CTA's half of stage st has landed =====
        if (!leader && elect_one()) {
            for (uint32_t it = 0; it < n_fill; ++it) {
                const int st = it % kStages;
                mbar_wait(&sh->full[st], (it / kStages) & 1);
                mbar_arrive_addr(mapa_u32(&sh->peer_full[st], 0));
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader only): one 256 x 48 x 16 MMA per K-step for the pair =====
        if (leader && elect_one()) {
            constexpr uint32_t idesc = make_idesc_f16(2 * kTileM, kChunkRows);
            uint32_t it = 0, step = 0, use[2] = {0, 0};
            for (int ip = 0; ip < n_p; ++ip)
                for (int t = 0; t < kSteps; ++t, ++step) {
                    mbar_wait(&sh->h_ready, step & 1);  // h_{t-1} complete in both CTAs' TMEM
                    tc_fence_after();
                    const uint32_t hold = tmem + ((t & 1) ? kColH1 : kColH0);
                    for (int c = 0; c < kChunks; ++c, ++it) {
                        const uint32_t g = c & 1;
                        mbar_wait(&sh->acc_empty[g], (use[g] & 1) ^ 1);
                        ++use[g];
                        const int st = it % kStages;
                        mbar_wait(&sh->full[st], (it / kStages) & 1);
                        mbar_wait(&sh->peer_full[st], (it / kStages) & 1);
                        tc_fence_after();
                        const uint32_t d = tmem + (g ? kColAcc1 : kColAcc0);
                        // K-step kk: h columns [8kk, 8kk+8); R slab 2 core columns = 2*24*16 B = 48 descriptor units on
                        const uint64_t bd0 = make_smem_desc(smem_u32(smem + st * kHalfBytes), kHalfRows * 16, 128);
                        const uint32_t blo = (uint32_t)bd0, bhi = (uint32_t)(bd0 >> 32);
                        mma2_ts_lohi<false>(d, hold, blo, bhi, idesc);
#pragma unroll
                        for (int kk = 1; kk < kH / 16; ++kk)
                            mma2_ts_lohi<true>(d, hold + kk * 8, blo + kk * (2 * kHalfRows), bhi, idesc);
                        mma2_commit_mc(&sh->empty[st], 3);
                        mma2_commit_mc(&sh->acc_full[g], 3);
                    }
                }
        }
    } else if (warp >= 4) {
        // ===== epilogue: group g = even / odd chunks; rows of THIS CTA's tile =====
        const int g = (warp - 4) >> 2;
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        const uint32_t lane_off = (uint32_t)(wq * 32) << 16;
        const uint32_t acc = tmem + lane_off + (g ? kColAcc1 : kColAcc0);
        const size_t row_off = (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16;
        uint32_t use = 0;
        for (int ip = 0; ip < n_p; ++ip) {
            const int q = 2 * (pair0 + ip * n_pair_ctas) + (int)rank;
            const bool valid = q < a.Q;  // an odd tile count leaves the last pair's second tile as padding
            // h_{-1} = 0: zero this group's columns of the ping buffer
            {
                uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                for (int c = g; c < kChunks; c += 2) tmem_st8(tmem + lane_off + kColH0 + c * 8, z);
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_addr(mapa_u32(&sh->h_ready, 0));
            }
            for (int t = 0; t < kSteps; ++t) {
                const uint32_t hold = tmem + lane_off + ((t & 1) ? kColH1 : kColH0);
                const uint32_t hnew = tmem + lane_off + ((t & 1) ? kColH0 : kColH1);
                const uint8_t *xpt = reinterpret_cast<const uint8_t *>(a.xp) + (size_t)(q * kSteps + t) * tile_bytes(3 * kH) + row_off;
                uint8_t *hot = reinterpret_cast<uint8_t *>(a.hout) + (size_t)(q * kSteps + t) * tile_bytes(kH) + row_off;
                const bool store = valid && t >= a.t_store_min;
                uint4 xn[6];
#pragma unroll
                for (int i = 0; i < 6; ++i)
                    xn[i] = valid ? __ldg(reinterpret_cast<const uint4 *>(xpt + (size_t)(g * 6 + i) * (kTileM * 16))) : make_uint4(0, 0, 0, 0);
                for (int c = g; c < kChunks; c += 2) {
                    uint4 x[6];
#pragma unroll
                    for (int i = 0; i < 6; ++i) x[i] = xn[i];
                    if (c + 2 < kChunks && valid) {
#pragma unroll
                        for (int i = 0; i < 6; ++i)
                            xn[i] = __ldg(reinterpret_cast<const uint4 *>(xpt + (size_t)((c + 2) * 6 + i) * (kTileM * 16)));
                    }
                    mbar_wait(&sh->acc_full[g], use & 1);
                    ++use;
                    tc_fence_after();
                    uint32_t az[16], ar[16], an[16], ho[8];
                    tmem_ld16(acc, az);
                    tmem_ld16(acc + 16, ar);
                    tmem_ld16(acc + 32, an);
                    tmem_ld8(hold + c * 8, ho);
                    tmem_wait_ld();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_addr(mapa_u32(&sh->acc_empty[g], 0));  // drained: the pair's next chunk may overwrite it

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
                if (lane == 0 && t < kSteps - 1) mbar_arrive_addr(mapa_u32(&sh->h_ready, 0));  // the next pair's zero fill arrives for its t = 0
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();  // neither CTA frees TMEM or exits while the other may still signal it
    if (warp == 2) tmem_dealloc2<512>(tmem);
}

}  // namespace

// Rimg2: per chunk [half 0 | half 1], each 24 gate rows x 400 K tiled with R = 24 (nn_tc.cu builds it)
void launch_tc_gru2(const GruArgs &a, int n_sm, cudaStream_t s) {
    static bool attr = false;
    if (!attr) {
        FVAD_CUDA_CHECK(cudaFuncSetAttribute(tc_gru2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGru2Smem));
        attr = true;
    }
    if (a.Q <= 0) return;
    const int n_pairs = (a.Q + 1) / 2;
    const int grid = 2 * (n_pairs < n_sm / 2 ? n_pairs : n_sm / 2);
    tc_gru2_kernel<<<grid, kGru2Threads, kGru2Smem, s>>>(a, n_pairs);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace tc
}  // namespace fvad
