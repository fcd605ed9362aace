// nn_tc_gru2.cu -- the GRU layer of nn_tc_gru.cu on CTA PAIRS (tcgen05 cta_group::2).
//
// Replaces the ONNX GRU nodes inside `onnx_instance.run()` (reference src/NSNet2.zig:220; ONNX GRU semantics with
// linear_before_reset = 1, gate order z,r,h -- SURVEY App. C); same math and data layout as nn_tc_gru.cu.
//
// What bounds a GRU step of 128 sequences on one SM (measured, tools/tmemld_bw.cu and the step timelines): 25 chunks x
// 25 MMAs of 128x48x16 = 15 000 tensor cycles, against (a) 0.96 MB of recurrent weights through L2 -> shared memory
// (64 B/clk/SM at that rate; the L2 delivers ~43 with every SM pulling) and (b) 25 x (48 accumulator + 8 hidden-state)
// columns x 128 lanes x 4 B = 717 KB through tcgen05.ld, which moves 48.5 B/clk/SM: 14 800 cycles.  Three pipes loaded
// to 100 % each cannot overlap perfectly; the one-CTA kernel runs at ~21 000 cycles per step.  This kernel takes load
// off (a) and (b):
//   * CTA pair: one 256-row MMA per K-step; each CTA keeps its own 128 sequences' hidden state and accumulators in its
//     own TMEM and only HALF of each R chunk (24 of the 48 gate rows) in its shared memory -- the weight stream per SM
//     halves.  Each half is a TMA tile load (.cta_group::2) whose bytes complete on the LEADER's barrier (no relay);
//     every barrier operation uses default CTA-scope semantics (tc_dev.cuh: mbar_arrive_addr);
//   * the shared memory the half-size ring frees (5 x 19 KB instead of 5 x 38 KB) holds h_{t-1} as fp16 in a
//     thread-private layout: the epilogue thread that wrote h_t[row][chunk] is the one that needs it for
//     h_{t+1} = n + z (h_t - n), so the blend operand comes from shared memory instead of a tcgen05.ld (-12 % drain);
//   * h-ready is signalled per GROUP OF 5 CHUNKS: K-step kk of step t+1 reads exactly the 16 hidden units chunk kk of
//     step t produced, so the issuers start step t+1 while the last epilogues of step t are still running;
//   * two MMA-issuing threads (even / odd chunks, one accumulator each) as in the one-CTA kernel.
//
// Roles (384 threads per CTA): warp 0 R producer (TMA), warps 1 and 3 MMA issuers (leader only; commits multicast to
// both CTAs), warp 2 TMEM allocator, warps 4..7 / 8..11 epilogue of the even / odd chunks (this CTA's 128 rows).
// TMEM columns: h ping [0,200) | h pong [200,400) | acc0 [400,448) | acc1 [448,496).
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
constexpr int kStages = 5;
constexpr int kGru2Threads = 384;                   // 4 control warps + 2 x 4 epilogue warps
constexpr int kHsBytes = kChunks * kTileM * 32;     // h_{t-1} as fp16, [chunk][row][16 units]: 102 400
constexpr int kGru2Smem = kStages * kHalfBytes + kHsBytes + kH * 4 + 1024;
constexpr uint32_t kColH0 = 0, kColH1 = 200, kColAcc0 = 400, kColAcc1 = 448;

// timeline probe: event e of (step t, chunk c) of the first tile pair of CTA 0 (the leader): 0 issuer acc_empty ok, 1 full ok,
// 2 issued + committed, 3 epilogue acc_full ok, 4 accumulator in registers, 5 chunk done (h signalled), 6 issuer starts chunk 0
#define GRU2_TRACE(e, t, c) do { if (a.trace && blockIdx.x == 0 && ip == 0) a.trace[((t) * 25 + (c)) * 8 + (e)] = (unsigned long long)clock64(); } while (0)

struct Gru2Shared {
    uint64_t full[kStages];       // LEADER's: both CTAs' halves of an R chunk complete their bytes here
    uint64_t empty[kStages];      // per CTA: the MMAs over this stage are done (multicast commit)
    uint64_t acc_full[2];         // per CTA (multicast commit)
    uint64_t acc_empty[2];        // LEADER's: 4 epilogue warps of the group in each CTA
    uint64_t h_rdy[kChunks / 5];  // LEADER's: the hidden units of chunks 5g..5g+4 of the step are in both CTAs' TMEM (5 x (4 + 4) warps)
    uint32_t tmem_base;
    uint32_t passed[2];           // LEADER's: per MMA issuer, 1 + the last chunk (producer's sequence) whose full-wait it has passed
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

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGru2Threads, 1)
    tc_gru2_kernel(const GruArgs a, int n_pairs_total, const __grid_constant__ CUtensorMap rmap) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *hs = smem + kStages * kHalfBytes;
    float *rbn_s = reinterpret_cast<float *>(hs + kHsBytes);
    Gru2Shared *sh = reinterpret_cast<Gru2Shared *>(hs + kHsBytes + kH * 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair0 = blockIdx.x >> 1, n_pair_ctas = gridDim.x >> 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&sh->full[i], 1);
            mbar_init(&sh->empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sh->acc_full[i], 1);
            mbar_init(&sh->acc_empty[i], 8);   // 4 epilogue warps of the group in each CTA
        }
        for (int i = 0; i < kChunks / 5; ++i) mbar_init(&sh->h_rdy[i], 40);
        sh->passed[0] = sh->passed[1] = 0;
        fence_mbar_init();
    }
    for (int i = threadIdx.x; i < kH; i += kGru2Threads) rbn_s[i] = a.rbn[i];
    if (warp == 2) tmem_alloc2<512>(&sh->tmem_base);
    tc_fence_before();
    __syncthreads();
    cluster_sync();  // both CTAs' barriers and TMEM exist before any remote arrive, remote complete_tx or paired MMA
    tc_fence_after();
    const uint32_t tmem = sh->tmem_base;

    const int n_p = (n_pairs_total - pair0 + n_pair_ctas - 1) / n_pair_ctas;  // tile pairs this cluster processes
    const uint32_t n_fill = (uint32_t)n_p * kSteps * kChunks;

    if (warp == 0) {
        // ===== producer: this CTA's half of every R chunk, identical sequence every step =====
        if (elect_one()) {
            const uint32_t full_leader = mapa_u32(&sh->full[0], 0);
            uint32_t it = 0, st = 0, ph = 0;
            for (int ip = 0; ip < n_p; ++ip)
                for (int t = 0; t < kSteps; ++t)
                    for (int c = 0; c < kChunks; ++c, ++it) {
                        mbar_wait(&sh->empty[st], ph ^ 1);
                        if (leader) mbar_expect_tx(&sh->full[st], 2 * kHalfBytes);
                        tma2_load_3d(smem + st * kHalfBytes, &rmap, full_leader + 8u * st, 0, 0, 2 * c + (int)rank);
                        if (++st == (uint32_t)kStages) { st = 0; ph ^= 1; }
                    }
            // drain: the leader's multicast commits to this CTA's `empty` barriers must have landed before exit
            for (uint32_t s = 0; s < (uint32_t)kStages && s < n_fill; ++s) {
                const uint32_t last = ((n_fill - 1 - s) / kStages) * kStages + s;  // last fill of stage s
                mbar_wait(&sh->empty[s], (last / kStages) & 1);
            }
        }
    } else if (warp == 1 || warp == 3) {
        // ===== MMA issuers (leader only): warp 1 the even chunks (accumulator 0), warp 3 the odd ones =====
        // One 256 x 48 x 16 MMA per K-step for the pair.  Chunk `it` of the producer's sequence sits in stage it % kStages;
        // the two threads skip each other's stages, and since a parity wait can only tell "this fill" from "the fill
        // before", a thread makes sure the other has seen the previous fill of a stage land before it polls that stage
        // (the previous fill is kStages = 5 chunks back: the other thread's whenever it is in the same step).
        if (leader && elect_one()) {
            constexpr uint32_t idesc = make_idesc_f16(2 * kTileM, kChunkRows);
            const uint32_t me = warp == 1 ? 0u : 1u;
            const uint32_t d = tmem + (me ? kColAcc1 : kColAcc0);
            volatile uint32_t *passed_me = &sh->passed[me], *passed_other = &sh->passed[me ^ 1u];
            uint32_t step = 0, use = 0;
            for (int ip = 0; ip < n_p; ++ip)
                for (int t = 0; t < kSteps; ++t, ++step) {
                    const uint32_t hold = tmem + ((t & 1) ? kColH1 : kColH0);
                    for (int c = (int)me; c < kChunks; c += 2) {
                        const uint32_t it = step * kChunks + (uint32_t)c;
                        const uint32_t st = it % kStages, ph = (it / kStages) & 1;
                        if (c == 0) GRU2_TRACE(6, t, 0);
                        // the weights first (they land long before), then the accumulator: the path drained -> first MMA is
                        // the critical one of the two-accumulator ring
                        if (it >= (uint32_t)kStages && (((it - kStages) % kChunks) & 1u) != me) {
                            while (*passed_other <= it - kStages) {}
                            __threadfence_block();
                        }
                        mbar_wait(&sh->full[st], ph);
                        __threadfence_block();
                        *passed_me = it + 1u;
                        GRU2_TRACE(1, t, c);
                        const uint64_t bd0 = make_smem_desc(smem_u32(smem + st * kHalfBytes), kHalfRows * 16, 128);
                        const uint32_t blo = (uint32_t)bd0, bhi = (uint32_t)(bd0 >> 32);
                        mbar_wait(&sh->acc_empty[me], (use & 1) ^ 1);
                        ++use;
                        tc_fence_after();
                        GRU2_TRACE(0, t, c);
                        // K-step kk: h columns [8kk, 8kk+8); R slab 2 core columns = 2*24*16 B = 48 descriptor units on
                        if (c < 2) {
                            // this thread's first chunk of the step: K-step kk needs the 16 hidden units that chunk kk of the
                            // previous step (or the zero fill) produced, in both CTAs; readiness comes in groups of 5 chunks
#pragma unroll
                            for (int grp = 0; grp < kChunks / 5; ++grp) {
                                mbar_wait(&sh->h_rdy[grp], step & 1);
                                tc_fence_after();
#pragma unroll
                                for (int kk = 5 * grp; kk < 5 * grp + 5; ++kk) {
                                    if (kk == 0) mma2_ts_lohi<false>(d, hold, blo, bhi, idesc);
                                    else mma2_ts_lohi<true>(d, hold + kk * 8, blo + kk * (2 * kHalfRows), bhi, idesc);
                                }
                            }
                        } else {
                            mma2_ts_lohi<false>(d, hold, blo, bhi, idesc);
#pragma unroll
                            for (int kk = 1; kk < kH / 16; ++kk)
                                mma2_ts_lohi<true>(d, hold + kk * 8, blo + kk * (2 * kHalfRows), bhi, idesc);
                        }
                        mma2_commit_mc(&sh->empty[st], 3);
                        mma2_commit_mc(&sh->acc_full[me], 3);
                        GRU2_TRACE(2, t, c);
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
        const uint32_t acc_empty_leader = mapa_u32(&sh->acc_empty[g], 0);
        const uint32_t h_rdy_leader = mapa_u32(&sh->h_rdy[0], 0);
        uint4 *hrow = reinterpret_cast<uint4 *>(hs + (size_t)row * 32);  // + chunk * (kTileM * 32) bytes
        uint32_t use = 0;
        for (int ip = 0; ip < n_p; ++ip) {
            const int q = 2 * (pair0 + ip * n_pair_ctas) + (int)rank;
            const bool valid = q < a.Q;  // an odd tile count leaves the last pair's second tile as padding
            // h_{-1} = 0: zero this group's columns of the ping buffer and of the shared-memory copy
            {
                uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                for (int c = g; c < kChunks; c += 2) {
                    tmem_st8(tmem + lane_off + kColH0 + c * 8, z);
                    hrow[c * (kTileM * 2)] = make_uint4(0, 0, 0, 0);
                    hrow[c * (kTileM * 2) + 1] = make_uint4(0, 0, 0, 0);
                }
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0)
                    for (int c = g; c < kChunks; c += 2) mbar_arrive_addr(h_rdy_leader + 8u * (uint32_t)(c / 5));
            }
            for (int t = 0; t < kSteps; ++t) {
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
                    const uint4 ho0 = hrow[c * (kTileM * 2)], ho1 = hrow[c * (kTileM * 2) + 1];  // h_{t-1} of this row's 16 units
                    mbar_wait(&sh->acc_full[g], use & 1);
                    ++use;
                    tc_fence_after();
                    if (wq == 0 && lane == 0) GRU2_TRACE(3, t, c);
                    uint32_t az[16], ar[16], an[16];
                    tmem_ld16(acc, az);
                    tmem_ld16(acc + 16, ar);
                    tmem_ld16(acc + 32, an);
                    tmem_wait_ld();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_addr(acc_empty_leader);  // drained: the pair's next chunk may overwrite it
                    if (wq == 0 && lane == 0) GRU2_TRACE(4, t, c);

                    const uint32_t *xw = reinterpret_cast<const uint32_t *>(x);  // 24 words: z[0..7] r[8..15] n[16..23] (pairs)
                    const uint32_t ho[8] = {ho0.x, ho0.y, ho0.z, ho0.w, ho1.x, ho1.y, ho1.z, ho1.w};
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
                    hrow[c * (kTileM * 2)] = make_uint4(hn[0], hn[1], hn[2], hn[3]);
                    hrow[c * (kTileM * 2) + 1] = make_uint4(hn[4], hn[5], hn[6], hn[7]);
                    if (store) {
                        uint8_t *p = hot + (size_t)(c * 2) * (kTileM * 16);
                        *reinterpret_cast<uint4 *>(p) = make_uint4(hn[0], hn[1], hn[2], hn[3]);
                        *reinterpret_cast<uint4 *>(p + kTileM * 16) = make_uint4(hn[4], hn[5], hn[6], hn[7]);
                    }
                    if (t < kSteps - 1) {  // h_t[16c..16c+15] of this warp's rows is in TMEM: K-step c of the next step may read it
                        tmem_wait_st();    // (the next pair's zero fill arrives for its own t = 0)
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_addr(h_rdy_leader + 8u * (uint32_t)(c / 5));
                    }
                    if (wq == 0 && lane == 0) GRU2_TRACE(5, t, c);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();  // neither CTA frees TMEM or exits while the other may still signal it
    if (warp == 2) tmem_dealloc2<512>(tmem);
}

}  // namespace

// Rimg2: per chunk [half 0 | half 1], each 24 gate rows x 400 K tiled with R = 24 (nn_tc.cu builds it): 50 half-chunk
// images of [50 core columns][24 rows x 16 B]; one box = one half-chunk.
void build_gru_pair_map(CUtensorMap &rmap, const void *Rimg2) {
    const cuuint64_t gdim[3] = {(cuuint64_t)kHalfRows * 2, (cuuint64_t)kH / 8, (cuuint64_t)2 * kChunks};  // 8-byte elements
    const cuuint64_t gstr[2] = {(cuuint64_t)kHalfRows * 16, (cuuint64_t)kHalfBytes};
    const cuuint32_t box[3] = {(cuuint32_t)kHalfRows * 2, (cuuint32_t)kH / 8, 1};
    encode_tiled_map(&rmap, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void *>(Rimg2), gdim, gstr, box);
}

void launch_tc_gru2(const GruArgs &a, const CUtensorMap &rmap, int n_sm, cudaStream_t s) {
    static bool attr = false;
    if (!attr) {
        FVAD_CUDA_CHECK(cudaFuncSetAttribute(tc_gru2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGru2Smem));
        attr = true;
    }
    if (a.Q <= 0) return;
    const int n_pairs = (a.Q + 1) / 2;
    const int grid = 2 * (n_pairs < n_sm / 2 ? n_pairs : n_sm / 2);
    tc_gru2_kernel<<<grid, kGru2Threads, kGru2Smem, s>>>(a, n_pairs, rmap);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace tc
}  // namespace fvad
