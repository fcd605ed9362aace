// nn_tc_gru2.cu -- the GRU layer of nn_tc_gru.cu on CTA PAIRS (tcgen05 cta_group::2).
//
// Replaces the ONNX GRU nodes inside `onnx_instance.run()` (reference src/NSNet2.zig:220; ONNX GRU semantics with
// linear_before_reset = 1, gate order z,r,h -- SURVEY App. C); same math and data layout as nn_tc_gru.cu.
//
// What bounds a GRU step of 128 sequences on one SM (step timelines, FVAD_GRU_TRACE; tools/mmabw_pair.cu,
// tools/tmemld_bw.cu): 25 chunks x 25 MMAs of 128x48x16 = 15 000 tensor cycles.  Neither the tensor pipe (pair MMAs
// run at N/2 cycles for every N) nor the accumulator drain (tcgen05.ld: 200 B/clk/SM with four warps) is the limit;
// the SM's L1 / shared-memory array is (128 B/clk): per chunk it takes the weight fill, the tensor cores' reads of
// it, the epilogue's x-projection loads (12 KB) and its h_t stores (8 KB).  One CTA: 64 + 64 + 33 = 161 B/clk at the
// tensor rate -> >= 19 000 cycles per step, ~21 000 measured.  This kernel:
//   * CTA pair: one 256-row MMA per K-step; each CTA keeps its own 128 sequences' hidden state and accumulators in its
//     own TMEM and only HALF of each R chunk (24 of the 48 gate rows) in its shared memory: fill 32 + reads 64 (own
//     tensor core + the peer's) B/clk.  Each half is a TMA tile load (.cta_group::2) whose bytes complete on the
//     LEADER's barrier (no relay); every barrier operation uses default CTA-scope semantics (tc_dev.cuh:
//     mbar_arrive_addr);
//   * the hidden state is SINGLE-buffered in TMEM (200 columns), which leaves room for SIX accumulators: the issuers
//     run up to six chunks ahead of the epilogue and the pipe never waits for a drain inside a step.  The epilogue
//     parks h_t in shared memory (fp16, thread-private layout -- the half-size ring frees the space) while the step's
//     MMAs still read h_{t-1} (which the blend h_t = n + z (h_{t-1} - n) also takes from TMEM); once the step's last
//     MMAs have retired each epilogue thread copies its chunks of h_t into TMEM for the next step;
//   * two MMA-issuing threads (even / odd chunks) as in the one-CTA kernel.
// Measured: ~20 500-21 400 cycles per step at a lower power draw (the step runs at the board's power cap, so the
// saved energy comes back as clock): 4 % faster than one CTA per 20 stream-hours.  The array is still the limit
// (96 + 46 B/clk); what would lift it is fusing the input projection into this kernel (no x-projection round trip),
// which needs TMEM for a second A operand.
//
// Roles (384 threads per CTA): warp 0 R producer (TMA), warps 1 and 3 MMA issuers (leader only; commits multicast to
// both CTAs), warp 2 TMEM allocator, warps 4..7 / 8..11 epilogue of the even / odd chunks (this CTA's 128 rows).
// TMEM columns: h [0,200) | accumulator a at [200 + 48a, 248 + 48a), a = (54-step chunk counter) % 6.
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
constexpr int kAccs = 6;                            // accumulators of 48 columns next to the 200 columns of h
constexpr uint32_t kColH = 0, kColAcc = 200;

// timeline probe: event e of (step t, chunk c) of the first tile pair of CTA 0 (the leader): 0 issuer acc_empty ok, 1 full ok,
// 2 issued + committed, 3 epilogue acc_full ok, 4 accumulator in registers, 5 chunk done (h signalled), 6 issuer starts chunk 0
#define GRU2_TRACE(e, t, c) do { if (a.trace && blockIdx.x == 0 && ip == 0) a.trace[((t) * 25 + (c)) * 8 + (e)] = (unsigned long long)clock64(); } while (0)

struct Gru2Shared {
    uint64_t full[kStages];       // LEADER's: both CTAs' halves of an R chunk complete their bytes here
    uint64_t empty[kStages];      // per CTA: the MMAs over this stage are done (multicast commit)
    uint64_t acc_full[kAccs];     // per CTA (multicast commit)
    uint64_t acc_empty[kAccs];    // LEADER's: 4 epilogue warps of the chunk's group in each CTA
    uint64_t step_done;           // per CTA (multicast commit by each issuer after its last chunk of a step): h may be overwritten
    uint64_t h_ready;             // LEADER's: h_t is in both CTAs' TMEM (8 + 8 epilogue warps)
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
// Gate arithmetic on packed FP32 pairs (FADD2 / FFMA2: one issue slot for two hidden units).  The arguments of the two
// sigmoids arrive HALVED (nn_tc.cu stores the z and r rows of W, R and their biases times 1/2), so
// sigma(x) = 0.5 + 0.5 tanh(x/2) is one packed add, two MUFU.TANH and one packed FMA per pair of units.
__device__ __forceinline__ float2 tanh2(float2 x) { return make_float2(tanh_hw(x.x), tanh_hw(x.y)); }
// experiments (FVAD_GRU_DBG 32 / 64): no MUFU at all (wrong results) / one MUFU.TANH.F16x2 for the pair (fp16 gate precision)
__device__ __forceinline__ float2 tanh2_none(float2 x) { return make_float2(x.x * 0.25f, x.y * 0.25f); }
__device__ __forceinline__ float2 tanh2_h2(float2 x) {
    uint32_t u = pack_h2(x.x, x.y), y;
    asm("tanh.approx.f16x2 %0, %1;" : "=r"(y) : "r"(u));
    return unpack_h2(y);
}

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
        for (int i = 0; i < kAccs; ++i) {
            mbar_init(&sh->acc_full[i], 1);
            mbar_init(&sh->acc_empty[i], 8);   // 4 epilogue warps of the group in each CTA
        }
        mbar_init(&sh->step_done, 2);
        mbar_init(&sh->h_ready, 16);
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
                        if (a.dbg & 2) { if (leader) mbar_arrive(&sh->full[st]); }
                        else {
                            if (leader) mbar_expect_tx(&sh->full[st], 2 * kHalfBytes);
                            tma2_load_3d(smem + st * kHalfBytes, &rmap, full_leader + 8u * st, 0, 0, 2 * c + (int)rank);
                        }
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
            const uint32_t hold = tmem + kColH;
            volatile uint32_t *passed_me = &sh->passed[me], *passed_other = &sh->passed[me ^ 1u];
            uint32_t step = 0;
            for (int ip = 0; ip < n_p; ++ip)
                for (int t = 0; t < kSteps; ++t, ++step) {
                    for (int c = (int)me; c < kChunks; c += 2) {
                        const uint32_t it = step * kChunks + (uint32_t)c;
                        const uint32_t st = it % kStages, ph = (it / kStages) & 1;
                        const uint32_t acc = it % kAccs, use = it / kAccs;
                        const uint32_t d = tmem + kColAcc + acc * kChunkRows;
                        if (c == 0) GRU2_TRACE(6, t, 0);
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
                        mbar_wait(&sh->acc_empty[acc], (use & 1) ^ 1);
                        if (c < 2) mbar_wait(&sh->h_ready, step & 1);  // h_{t-1} (or the zero fill) is in both CTAs' TMEM
                        tc_fence_after();
                        GRU2_TRACE(0, t, c);
                        // K-step kk: h columns [8kk, 8kk+8); R slab 2 core columns = 2*24*16 B = 48 descriptor units on
                        mma2_ts_lohi<false>(d, hold, blo, bhi, idesc);
#pragma unroll
                        for (int kk = 1; kk < kH / 16; ++kk)
                            mma2_ts_lohi<true>(d, hold + kk * 8, blo + kk * (2 * kHalfRows), bhi, idesc);
                        mma2_commit_mc(&sh->empty[st], 3);
                        mma2_commit_mc(&sh->acc_full[acc], 3);
                        GRU2_TRACE(2, t, c);
                    }
                    mma2_commit_mc(&sh->step_done, 3);  // every MMA of this thread that reads h_{t-1} has retired when this lands
                }
        }
    } else if (warp >= 4) {
        // ===== epilogue: group g = even / odd chunks; rows of THIS CTA's tile =====
        const int g = (warp - 4) >> 2;
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        const uint32_t lane_off = (uint32_t)(wq * 32) << 16;
        const size_t row_off = (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16;
        const uint32_t acc_empty_leader = mapa_u32(&sh->acc_empty[0], 0);
        const uint32_t h_ready_leader = mapa_u32(&sh->h_ready, 0);
        const uint32_t th = tmem + lane_off + kColH;
        uint4 *hrow = reinterpret_cast<uint4 *>(hs + (size_t)row * 32);  // + chunk * (kTileM * 32) bytes
        uint32_t step = 0;
        for (int ip = 0; ip < n_p; ++ip) {
            const int q = 2 * (pair0 + ip * n_pair_ctas) + (int)rank;
            const bool valid = q < a.Q;  // an odd tile count leaves the last pair's second tile as padding
            // h_{-1} = 0: zero this group's chunks of h in TMEM (the previous tile's last MMAs have retired: the step_done
            // wait below)
            {
                uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                for (int c = g; c < kChunks; c += 2) {
                    tmem_st8(th + c * 8, z);
                }
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_addr(h_ready_leader);
            }
            for (int t = 0; t < kSteps; ++t, ++step) {
                const uint8_t *xpt = reinterpret_cast<const uint8_t *>(a.xp) + (size_t)(q * kSteps + t) * tile_bytes(3 * kH) + row_off;
                uint8_t *hot = reinterpret_cast<uint8_t *>(a.hout) + (size_t)(q * kSteps + t) * tile_bytes(kH) + row_off;
                const bool store = valid && t >= a.t_store_min;
                uint4 xn[6];
#pragma unroll
                for (int i = 0; i < 6; ++i)
                    xn[i] = (valid && !(a.dbg & 4)) ? ldg_stream(xpt + (size_t)(g * 6 + i) * (kTileM * 16)) : make_uint4(0, 0, 0, 0);
                for (int c = g; c < kChunks; c += 2) {
                    uint4 x[6];
#pragma unroll
                    for (int i = 0; i < 6; ++i) x[i] = xn[i];
                    if (c + 2 < kChunks && valid && !(a.dbg & 4)) {
#pragma unroll
                        for (int i = 0; i < 6; ++i)
                            xn[i] = ldg_stream(xpt + (size_t)((c + 2) * 6 + i) * (kTileM * 16));
                    }
                    const uint32_t it = step * kChunks + (uint32_t)c;
                    const uint32_t ai = it % kAccs, use = it / kAccs;
                    const uint32_t acc = tmem + lane_off + kColAcc + ai * kChunkRows;
                    mbar_wait(&sh->acc_full[ai], use & 1);
                    tc_fence_after();
                    if (wq == 0 && lane == 0) GRU2_TRACE(3, t, c);
                    uint32_t az[16], ar[16], an[16], ho[8];
                    if (a.dbg & 1) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) { az[i] = ar[i] = an[i] = 0; ho[i & 7] = 0; }
                    } else {
                        tmem_ld16(acc, az);
                        tmem_ld16(acc + 16, ar);
                        tmem_ld16(acc + 32, an);
                        tmem_ld8(th + c * 8, ho);  // h_{t-1} of this row's 16 units: TMEM keeps it untouched for the whole step
                    }
                    tmem_wait_ld();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_addr(acc_empty_leader + 8u * ai);  // drained: a later chunk may overwrite it
                    if (wq == 0 && lane == 0) GRU2_TRACE(4, t, c);

                    const uint32_t *xw = reinterpret_cast<const uint32_t *>(x);  // 24 words: z[0..7] r[8..15] n[16..23] (pairs)
                    // Rb of the n gate for the chunk's 16 units: four broadcast 16-byte loads (the L1 / shared-memory pipe is the
                    // loaded resource of this kernel: weights in and out, the x projections and h_t all go through it)
                    float rb[16];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 v = *reinterpret_cast<const float4 *>(rbn_s + c * kChunkUnits + 4 * i);
                        rb[4 * i] = v.x; rb[4 * i + 1] = v.y; rb[4 * i + 2] = v.z; rb[4 * i + 3] = v.w;
                    }
                    uint32_t hn[8];
                    if (a.dbg & 16) {
#pragma unroll
                        for (int p = 0; p < 8; ++p) hn[p] = xw[p] ^ az[p] ^ ho[p];
                    } else {
                        const float2 kHalf = make_float2(0.5f, 0.5f), kNeg = make_float2(-1.0f, -1.0f);
                        auto gates = [&](auto th) {
#pragma unroll
                            for (int p = 0; p < 8; ++p) {
                                const float2 xz = unpack_h2(xw[p]), xr = unpack_h2(xw[8 + p]), xnn = unpack_h2(xw[16 + p]);
                                const float2 hp = unpack_h2(ho[p]);
                                const float2 az2 = make_float2(__uint_as_float(az[2 * p]), __uint_as_float(az[2 * p + 1]));
                                const float2 ar2 = make_float2(__uint_as_float(ar[2 * p]), __uint_as_float(ar[2 * p + 1]));
                                const float2 an2 = make_float2(__uint_as_float(an[2 * p]), __uint_as_float(an[2 * p + 1]));
                                const float2 b2 = make_float2(rb[2 * p], rb[2 * p + 1]);
                                const float2 z = __ffma2_rn(th(__fadd2_rn(xz, az2)), kHalf, kHalf);
                                const float2 r = __ffma2_rn(th(__fadd2_rn(xr, ar2)), kHalf, kHalf);
                                const float2 n = th(__ffma2_rn(r, __fadd2_rn(an2, b2), xnn));
                                const float2 h = __ffma2_rn(z, __ffma2_rn(n, kNeg, hp), n);  // n + z (h_{t-1} - n)
                                hn[p] = pack_h2(h.x, h.y);
                            }
                        };
                        if (a.dbg & 32) gates([](float2 x) { return tanh2_none(x); });
                        else if (a.dbg & 64) gates([](float2 x) { return tanh2_h2(x); });
                        else gates([](float2 x) { return tanh2(x); });
                    }
                    hrow[c * (kTileM * 2)] = make_uint4(hn[0], hn[1], hn[2], hn[3]);
                    hrow[c * (kTileM * 2) + 1] = make_uint4(hn[4], hn[5], hn[6], hn[7]);
                    if (store && !(a.dbg & 8)) {
                        uint8_t *p = hot + (size_t)(c * 2) * (kTileM * 16);
                        *reinterpret_cast<uint4 *>(p) = make_uint4(hn[0], hn[1], hn[2], hn[3]);
                        *reinterpret_cast<uint4 *>(p + kTileM * 16) = make_uint4(hn[4], hn[5], hn[6], hn[7]);
                    }
                    if (wq == 0 && lane == 0) GRU2_TRACE(5, t, c);
                }
                // The step's last MMAs (both issuers) have read h_{t-1}: its TMEM columns may now take h_t, which this
                // thread holds in shared memory for the chunks it computed.  (After the tile's last step the next tile's
                // zero fill follows instead.)
                mbar_wait(&sh->step_done, step & 1);
                tc_fence_after();
                if (t < kSteps - 1) {
                    for (int c = g; c < kChunks; c += 2) {
                        const uint4 v0 = hrow[c * (kTileM * 2)], v1 = hrow[c * (kTileM * 2) + 1];
                        const uint32_t hv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
                        tmem_st8(th + c * 8, hv);
                    }
                    tmem_wait_st();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_addr(h_ready_leader);
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
