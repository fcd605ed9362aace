// nn_tc_gemm2.cu -- the dense-layer GEMM of nn_tc_gemm.cu on CTA PAIRS (tcgen05 cta_group::2, M = 256), with the
// weight halves delivered by TMA straight onto the MMA leader's barrier.
//
// Replaces the MatMul/Add/Relu/Sigmoid nodes that onnxruntime executes inside `onnx_instance.run()` (reference
// src/NSNet2.zig:220) -- fc1, the GRU input projections W x + b for all 54 steps at once, fc2, fc3, fc4 -- like
// nn_tc_gemm.cu, whose math, operand layouts and epilogue it shares.
//
// Why pairs.  A one-CTA M = 128 tile needs 2 bytes of weights per 2*128 flop: 64 B/clk/SM at the full tensor rate,
// next to ~27 B/clk/SM of activation reads and writes, against an L2 that delivers ~43 B/clk/SM with all SMs pulling
// (B300_MICROARCH: ~6 300 B/clk chip-wide, path-independent; measured here as 58 B/clk at the power-capped SM
// clock, profiles/r01_gemm_timeline_a_in_tmem.txt).  The one-CTA kernel is therefore L2-bound at ~2/3 of the tensor
// rate whatever its pipeline depth.  In a pair each CTA stores HALF of every weight stage (N/2 rows) and the tensor
// cores of both SMs read both halves, so the weight stream per SM halves.
//
// What makes it fast (the first pair kernel, with a relay thread forwarding the peer's barrier completions and
// cluster-scope barrier operations, was 50 % slower than one CTA):
//   * each CTA's half of a weight stage is a TMA tile load (cp.async.bulk.tensor, .cta_group::2) whose bytes complete
//     on the LEADER's `full` barrier -- the MMA issuer waits on one local barrier for both halves, no relay;
//   * every barrier arrive and wait, the remote ones too, uses the default CTA-scope semantics (tc_dev.cuh:
//     mbar_arrive_addr): `.acquire.cluster` polls stalled the issuer until its own MMAs had retired and
//     `.release.cluster` arrives made the epilogue drain its global stores before handing an accumulator back;
//   * the A tile (this CTA's own 128 rows) lives in TENSOR MEMORY (K/2 columns, A-from-TMEM MMAs): its K slabs come
//     through their own staging buffers with their own producer thread, so they never occupy weight stages (in the
//     one-CTA kernel the next tile's A slabs sit in the ring until the previous tile's MMAs free the TMEM columns,
//     which stalls the weight prefetch for a whole L2 latency at every M-tile change);
//   * two MMA-issuing threads, one per accumulator (even / odd N-tiles): one's barrier round trips, descriptor
//     arithmetic and commits hide under the other's issue burst (an N = 96 slab of 8 MMAs is 384 tensor cycles,
//     a lone issuer needs ~450 per slab);
//   * EIGHT epilogue warps (two sets on alternate 32-column chunks): the epilogue is bound by the dependent-issue
//     latency of one warp per scheduler (~30 instructions per 8 columns: bias, activation, fp16 pack, 16-byte store),
//     not by tcgen05.ld (200 B/clk/SM with four warps, 357 with eight: tools/tmemld_bw.cu), so a second warp per
//     scheduler nearly halves it; the accumulator goes back to the issuer as soon as a warp's last chunk sits in registers.
//
// Roles (both CTAs run all of them except the MMA issuers): warp 0 weight producer (TMA), warp 3 A producer
// (cp.async.bulk), warps 1 and 2 MMA issuers (leader only; commits multicast to both CTAs; warp 2 also owns the TMEM
// allocation), warps 4..7 and 12..15 epilogue (own 128 rows, alternate 32-column chunks), warps 8..11 A stagers
// (staging buffer -> TMEM, tcgen05.st).  512 threads x <= 112 registers (__maxnreg__): the VAD scan's one-warp CTAs must
// still fit on the SM (a GEMM CTA that owns the whole register file cannot start next to them: measured, fc3 +60 %).
// NSETS = 1 (384 threads, loads software-pipelined one chunk ahead) is kept as a template option.
#include <cuda.h>

#include <cstdlib>

#include "nn_tc.cuh"
#include "tc_dev.cuh"

namespace fvad {
namespace tc {

namespace {

constexpr int kMaxSlabs = 8;
constexpr int kMaxStages = 16;
constexpr int kMaxABuf = 5;
constexpr int kSmemTotal = 214 * 1024;  // see nn_tc_gemm.cu: leaves room for the concurrently running VAD scan

// timeline probe (CTA 0 = the leader of pair 0, first 4 M-pairs, first 8 N-tiles): [kind][mpair][ntile][slot], kinds:
// 0 weight producer issued slab, 1 MMA full ok, 2 MMA slab issued+committed, 3 epilogue acc_full ok (slot 0) / done (slot 1)
// and MMA a_full[sl] ok (slot 2+sl, N-tile 0), 4 MMA acc_empty ok (slot 0) / accumulator committed (slot 1) and the
// stagers' a_empty ok / as_full ok / staged / stored (slot 2+sl of N-tiles 0 / 1 / 2 / 3)
#define PAIR_TRACE(kind, im, j, sl) do { if (a.trace && blockIdx.x == 0 && (im) < 4 && (j) < 8) \
    a.trace[(((kind) * 4 + (im)) * 8 + (j)) * 32 + (sl)] = (unsigned long long)clock64(); } while (0)

struct PairShared {
    uint64_t full[kMaxStages];    // LEADER's copy is the live one: both CTAs' weight halves complete their bytes here
    uint64_t empty[kMaxStages];   // per CTA: the leader's MMAs over this stage are done (multicast commit)
    uint64_t as_full[kMaxABuf];   // per CTA: an A slab has landed in staging buffer i
    uint64_t as_empty[kMaxABuf];  // per CTA: the stagers have moved it to TMEM
    uint64_t a_full[kMaxSlabs];   // LEADER's: A slab sl is in TMEM in BOTH CTAs (2 arrivals)
    uint64_t a_empty[kMaxSlabs];  // per CTA: the MMAs of the M-pair's last two N-tiles (one per issuer) over TMEM slab sl are done
    uint64_t acc_full[3];         // per CTA: accumulator complete (multicast commit)
    uint64_t acc_empty[3];        // LEADER's: drained by the epilogue warps of both CTAs (8 * NSETS arrivals)
    uint32_t tmem_base;
    uint32_t passed[2];           // LEADER's: per MMA issuer, 1 + the last slab (producer's sequence) whose full-wait it has passed
};

struct PairPlan {  // host-computed (launch_tc_gemm2)
    int ks;             // K-steps of 16 columns per slab
    int nk_total;       // K / 16
    int n_slabs, n_stages, n_abuf;
    int stage_bytes;    // per CTA: max over N-tiles of (NT/2 rows) x (ks K-steps) fp16
    int a_slab_bytes;   // 128 rows x ks K-steps fp16
    int b_off, bias_off, tr_off, sh_off, smem_bytes;
    int acc_cols;       // TMEM columns between accumulators (max NT)
    int n_acc;          // accumulators: 3 when 3*acc_cols + K/2 <= 512 (N-tile `work` uses accumulator work % n_acc), else 2; A follows them
};

// GATHER (K = 176, two slabs of 6 + 5 K-steps): the A operand is gathered from the fp16 feature rows by the stager warps
// (thread = row: 22 16-byte loads per row, issued for the NEXT M-tile as soon as this one is in TMEM, so they fly under a whole
// M-tile of MMAs), see GemmArgs::gather_rows.  No staging buffers, no A producer: all of shared memory is weight ring.
constexpr int kGatherKsteps = 11, kGatherKs = 6;

template <int ACT, int OUT, int NSETS, bool GATHER>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(384 + 128 * (NSETS - 1), 1) __maxnreg__(NSETS == 2 ? 112 : 128)
    tc_gemm_pair_kernel(const GemmArgs a, const PairPlan pl, const __grid_constant__ PairMaps maps) {
    constexpr int kThreads = 384 + 128 * (NSETS - 1);
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sAst = smem, *sB = smem + pl.b_off;
    float *bias_s = reinterpret_cast<float *>(smem + pl.bias_off);
    PairShared *sh = reinterpret_cast<PairShared *>(smem + pl.sh_off);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int pair0 = blockIdx.x >> 1, pair_stride = gridDim.x >> 1;
    const int n_mpairs = (a.mt_count + 1) >> 1;
    const int ks = pl.ks, n_slabs = pl.n_slabs, n_stages = pl.n_stages, n_abuf = pl.n_abuf;

    if (threadIdx.x == 0) {
        for (int i = 0; i < n_stages; ++i) {
            mbar_init(&sh->full[i], 1);
            mbar_init(&sh->empty[i], 1);
        }
        for (int i = 0; i < n_abuf; ++i) {
            mbar_init(&sh->as_full[i], 1);
            mbar_init(&sh->as_empty[i], 1);
        }
        for (int i = 0; i < n_slabs; ++i) {
            mbar_init(&sh->a_full[i], 2);
            mbar_init(&sh->a_empty[i], a.n_tiles >= 2 ? 2 : 1);
        }
        for (int i = 0; i < pl.n_acc; ++i) {
            mbar_init(&sh->acc_full[i], 1);
            mbar_init(&sh->acc_empty[i], 8 * NSETS);
        }
        sh->passed[0] = sh->passed[1] = 0;
        fence_mbar_init();
    }
    for (int i = threadIdx.x; i < a.Nc; i += kThreads) bias_s[i] = a.bias[i];
    if (warp == 2) tmem_alloc2<512>(&sh->tmem_base);
    tc_fence_before();
    __syncthreads();
    cluster_sync();  // both CTAs' barriers and TMEM exist before any remote arrive, remote complete_tx or paired MMA
    tc_fence_after();
    const uint32_t tmem = sh->tmem_base;
    const size_t a_tile_bytes = tile_bytes(a.K);
    const uint32_t tA0 = tmem + (uint32_t)(pl.n_acc * pl.acc_cols);  // A image: 8 columns per K-step
    const uint32_t n_acc = (uint32_t)pl.n_acc;

    if (warp == 0) {
        // ===== weight producer: this CTA's half (NT/2 rows) of every stage, bytes completed on the leader's barrier =====
        if (elect_one()) {
            uint32_t it = 0, st = 0, ph = 0, im = 0;
            const uint32_t full_leader = mapa_u32(&sh->full[0], 0);
            for (int mp = pair0; mp < n_mpairs; mp += pair_stride, ++im) {
                for (int j = 0; j < a.n_tiles; ++j) {
                    const int NH = a.nt_size[j] >> 1;
                    const void *tmap = &maps.m[2 * j + (int)rank];
                    const uint32_t both = 2u * (uint32_t)(NH * 16) * (uint32_t)(2 * ks);  // the box is always whole (columns past K are zero-filled)
                    for (int sl = 0; sl < n_slabs; ++sl, ++it) {
                        mbar_wait(&sh->empty[st], ph ^ 1);
                        if (a.dbg & 2) { if (leader) mbar_arrive(&sh->full[st]); }  // (experiments: no weight copies)
                        else {
                            if (leader) mbar_expect_tx(&sh->full[st], both);
                            tma2_load_2d(sB + (size_t)st * pl.stage_bytes, tmap, full_leader + 8u * st, 0, sl * 2 * ks);
                        }
                        PAIR_TRACE(0, im, j, sl);
                        if (++st == (uint32_t)n_stages) { st = 0; ph ^= 1; }
                    }
                }
            }
            // drain: the leader's multicast commits to this CTA's barriers must have landed before the CTA exits
            for (uint32_t s = 0; s < (uint32_t)n_stages && s < it; ++s) {
                const uint32_t last = ((it - 1 - s) / n_stages) * n_stages + s;
                mbar_wait(&sh->empty[s], (last / n_stages) & 1);
            }
        }
    } else if (warp == 3) {
        // ===== A producer: this CTA's own M-tile, slab by slab into the staging buffers =====
        if (!GATHER && elect_one()) {
            uint32_t ab = 0, abph = 0;
            for (int mp = pair0; mp < n_mpairs; mp += pair_stride) {
                const int mt = 2 * mp + (int)rank;
                if (mt >= a.mt_count) continue;  // an odd tile count leaves the last pair's second tile as padding
                const int q = mt / a.T, t = mt % a.T;
                const uint8_t *At = reinterpret_cast<const uint8_t *>(a.A) + (size_t)(q * a.A_q + t + a.A_t0) * a_tile_bytes;
                for (int sl = 0; sl < n_slabs; ++sl) {
                    const int nk = min(ks, pl.nk_total - sl * ks);
                    mbar_wait(&sh->as_empty[ab], abph ^ 1);
                    const uint32_t bytes = (uint32_t)nk * (2u * kTileM * 16u);
                    mbar_expect_tx(&sh->as_full[ab], bytes);
                    bulk_g2s(sAst + (size_t)ab * pl.a_slab_bytes, At + (size_t)(sl * ks) * (2 * kTileM * 16), bytes, &sh->as_full[ab]);
                    if (++ab == (uint32_t)n_abuf) { ab = 0; abph ^= 1; }
                }
            }
        }
    } else if (warp == 1 || warp == 2) {
        // ===== MMA issuers (leader only): warp 1 takes the even N-tiles, warp 2 the odd ones =====
        // 256 x NT x 16 per instruction, A from each CTA's own TMEM.  N-tile `work` accumulates into buffer work % n_acc
        // and uses ring stages [work*n_slabs, (work+1)*n_slabs) of the producer's sequence, so the two threads never
        // share an accumulator or a stage.
        if (leader && elect_one()) {
            const uint32_t me = warp == 1 ? 0u : 1u;
            // One ring, consumed by two threads that skip each other's stages.  A parity wait can only tell "this fill" from
            // "the fill before", so before polling a stage whose previous fill belonged to the other issuer's N-tile this
            // thread makes sure the other one has seen that fill land (sh->passed[]: slabs whose full-wait it has passed).
            // The previous fill of slab (work, sl) is n_stages slabs back: N-tile work - q - (sl < rr).
            const uint32_t q = (uint32_t)(n_stages / n_slabs), rr = (uint32_t)(n_stages % n_slabs);
            volatile uint32_t *passed_me = &sh->passed[me], *passed_other = &sh->passed[me ^ 1u];
            uint32_t st = 0, ph = 0, work = 0, im = 0;
            for (int mp = pair0; mp < n_mpairs; mp += pair_stride, ++im) {
                for (int j = 0; j < a.n_tiles; ++j, ++work) {
                    if ((work & 1u) != me) {  // the other issuer's N-tile: skip its stages
                        st += (uint32_t)n_slabs;
                        while (st >= (uint32_t)n_stages) { st -= (uint32_t)n_stages; ph ^= 1; }
                        continue;
                    }
                    const int NT = a.nt_size[j], NH = NT >> 1;
                    const uint32_t buf = work % n_acc, use = work / n_acc;
                    const uint32_t d = tmem + buf * (uint32_t)pl.acc_cols;
                    mbar_wait(&sh->acc_empty[buf], (use & 1) ^ 1);
                    tc_fence_after();
                    PAIR_TRACE(4, im, j, 0);
                    const uint32_t idesc = make_idesc_f16(2 * kTileM, NT);
                    const bool first_mine = j <= 1;                 // this thread's first N-tile of the M-pair: the A slabs may still be arriving
                    const bool frees_a = j >= a.n_tiles - 2;        // the M-pair's last N-tile of each issuer releases the A slabs
                    const uint32_t bstep = 2u * (uint32_t)NH;       // one K-step of a half stage = 2 core columns = 2*NH*16 bytes
                    for (int sl = 0; sl < n_slabs; ++sl) {
                        if (first_mine) { mbar_wait(&sh->a_full[sl], im & 1); if (j == 0) PAIR_TRACE(3, im, 0, 2 + sl); }
                        {
                            const uint32_t back = q + ((uint32_t)sl < rr ? 1u : 0u);  // N-tiles between this slab and its stage's previous fill
                            const uint32_t idx = work * (uint32_t)n_slabs + (uint32_t)sl;  // this slab in the producer's sequence
                            if ((back & 1u) && idx >= (uint32_t)n_stages) {
                                while (*passed_other <= idx - (uint32_t)n_stages) {}
                                __threadfence_block();
                            }
                        }
                        mbar_wait(&sh->full[st], ph);
                        __threadfence_block();
                        *passed_me = work * (uint32_t)n_slabs + (uint32_t)sl + 1u;
                        tc_fence_after();
                        PAIR_TRACE(1, im, j, sl);
                        const int nk = min(ks, pl.nk_total - sl * ks);
                        const uint64_t bd0 = make_smem_desc(smem_u32(sB + (size_t)st * pl.stage_bytes), NH * 16, 128);
                        const uint32_t blo = (uint32_t)bd0, bhi = (uint32_t)(bd0 >> 32);
                        const uint32_t ta = tA0 + (uint32_t)(sl * ks) * 8u;
                        if (sl == 0) mma2_ts_lohi<false>(d, ta, blo, bhi, idesc);
                        else mma2_ts_lohi<true>(d, ta, blo, bhi, idesc);
#pragma unroll
                        for (int kk = 1; kk < 8; ++kk)
                            if (kk < nk) mma2_ts_lohi<true>(d, ta + kk * 8u, blo + kk * bstep, bhi, idesc);
                        mma2_commit_mc(&sh->empty[st], 3);                 // the stage is free in both CTAs once these MMAs have read it
                        if (frees_a) mma2_commit_mc(&sh->a_empty[sl], 3);  // and (with the other issuer's commit) so are this A slab's TMEM columns
                        PAIR_TRACE(2, im, j, sl);
                        if (++st == (uint32_t)n_stages) { st = 0; ph ^= 1; }
                    }
                    mma2_commit_mc(&sh->acc_full[buf], 3);
                    PAIR_TRACE(4, im, j, 1);
                }
            }
        }
    } else if (warp >= 8 && warp < 12) {
        // ===== A stagers: staging buffer -> TMEM, thread = row (lane quadrant = warp % 4) =====
        const int row = (warp & 3) * 32 + lane;
        const uint32_t tA = tA0 + ((uint32_t)((warp & 3) * 32) << 16);
        const uint32_t a_full_leader = mapa_u32(&sh->a_full[0], 0);
        uint32_t ab = 0, abph = 0, im = 0;
        if (GATHER) {
            uint4 d[2 * kGatherKsteps];  // this thread's row of the current M-tile: 11 K-steps x 32 bytes
            auto fetch = [&](int mp) {
                const int mt = 2 * mp + (int)rank;
                const int q = mt / a.T, t = mt % a.T;
                const int g = a.seq0 + q * kTileM + row;
                const bool ok = mt < a.mt_count && g < a.n_seq;
                const int sc = ok ? g / a.Kchunks : 0, j = ok ? g - sc * a.Kchunks : 0;
                const uint4 *src = reinterpret_cast<const uint4 *>(a.gather_rows + ((size_t)sc * (kCtx + kFrames * a.Kchunks) + (size_t)kFrames * j + t + a.A_t0) * 176);
#pragma unroll
                for (int i = 0; i < 2 * kGatherKsteps; ++i) d[i] = ok ? __ldg(src + i) : make_uint4(0u, 0u, 0u, 0u);
            };
            if (pair0 < n_mpairs) fetch(pair0);
            for (int mp = pair0; mp < n_mpairs; mp += pair_stride, ++im) {
#pragma unroll
                for (int sl = 0; sl < 2; ++sl) {
                    mbar_wait(&sh->a_empty[sl], (im & 1) ^ 1);  // the previous M-tile's MMAs over these columns are done
                    tc_fence_after();
#pragma unroll
                    for (int kk = 0; kk < kGatherKs; ++kk) {
                        const int ks_abs = sl * kGatherKs + kk;
                        if (ks_abs < kGatherKsteps) {
                            const uint32_t r[8] = {d[2 * ks_abs].x, d[2 * ks_abs].y, d[2 * ks_abs].z, d[2 * ks_abs].w,
                                                   d[2 * ks_abs + 1].x, d[2 * ks_abs + 1].y, d[2 * ks_abs + 1].z, d[2 * ks_abs + 1].w};
                            tmem_st8(tA + (uint32_t)ks_abs * 8u, r);
                        }
                    }
                    tmem_wait_st();
                    tc_fence_before();
                    bar_sync(1, 128);
                    if (warp == 8 && lane == 0) mbar_arrive_addr(a_full_leader + 8u * (uint32_t)sl);
                }
                if (mp + pair_stride < n_mpairs) fetch(mp + pair_stride);
            }
        }
        for (int mp = pair0; !GATHER && mp < n_mpairs; mp += pair_stride, ++im) {
            const bool valid = 2 * mp + (int)rank < a.mt_count;
            for (int sl = 0; sl < n_slabs; ++sl) {
                const int nk = min(ks, pl.nk_total - sl * ks);
                mbar_wait(&sh->a_empty[sl], (im & 1) ^ 1);  // the previous M-tile's MMAs over these columns are done
                tc_fence_after();
                if (warp == 8 && lane == 0) PAIR_TRACE(4, im, 0, 2 + sl);
                if (valid) {
                    mbar_wait(&sh->as_full[ab], abph);
                    if (warp == 8 && lane == 0) PAIR_TRACE(4, im, 1, 2 + sl);
                    const uint8_t *src = sAst + (size_t)ab * pl.a_slab_bytes + (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16;
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        if (kk < nk) {
                            const uint4 lo = *reinterpret_cast<const uint4 *>(src + (size_t)(2 * kk) * (kTileM * 16));
                            const uint4 hi = *reinterpret_cast<const uint4 *>(src + (size_t)(2 * kk + 1) * (kTileM * 16));
                            const uint32_t r[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
                            tmem_st8(tA + (uint32_t)(sl * ks + kk) * 8u, r);
                        }
                    }
                    tmem_wait_st();
                }
                if (warp == 8 && lane == 0) PAIR_TRACE(4, im, 3, 2 + sl);
                tc_fence_before();
                bar_sync(1, 128);
                if (warp == 8 && lane == 0) {
                    if (valid) mbar_arrive(&sh->as_empty[ab]);
                    mbar_arrive_addr(a_full_leader + 8u * (uint32_t)sl);
                    PAIR_TRACE(4, im, 2, 2 + sl);
                }
                if (valid && ++ab == (uint32_t)n_abuf) { ab = 0; abph ^= 1; }
            }
        }
        // drain the leader's last multicast commits to this CTA's a_empty barriers
        if (im) for (int sl = 0; sl < n_slabs; ++sl) mbar_wait(&sh->a_empty[sl], (im - 1) & 1);
    } else if (warp >= 4) {
        // ===== epilogue (this CTA's 128 rows): warp set h takes the 32-column chunks with (c/32) % NSETS == h =====
        const int wq = warp & 3;
        const int hset = warp >= 12 ? 1 : 0;
        const int row = wq * 32 + lane;
        const size_t c_tile_bytes = tile_bytes(a.Nc);
        float *tr = reinterpret_cast<float *>(smem + pl.tr_off) + (hset * 4 + wq) * (32 * 33);  // OUT == 1: this warp's 32x32 transpose pad
        uint32_t work = 0, im = 0;
        const uint32_t acc_empty_leader = mapa_u32(&sh->acc_empty[0], 0);
        uint8_t *Ct = nullptr;
        float *g0 = nullptr;   // OUT == 1: gains row of this warp's first sequence
        int rows_valid = 0;

        // bias + activation of 8 consecutive columns starting at column `col` (absolute, multiple of 8)
        auto act8 = [&](const uint32_t *r, int col, float (&v)[8]) {
            const float4 b0 = *reinterpret_cast<const float4 *>(bias_s + col), b1 = *reinterpret_cast<const float4 *>(bias_s + col + 4);
            const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float x = __uint_as_float(r[i]) + b[i];
                if (ACT == 1) x = fmaxf(x, 0.0f);
                if (ACT == 2) x = __fdividef(1.0f, 1.0f + __expf(-x));
                v[i] = x;
            }
        };
        // one chunk of up to 32 accumulator columns (ng groups of 8), absolute first column `col`
        auto emit = [&](const uint32_t (&r)[32], int col, int ng) {
            if (OUT == 0) {
#pragma unroll
                for (int gI = 0; gI < 4; ++gI) {
                    if (gI < ng) {
                        float v[8];
                        act8(r + 8 * gI, col + 8 * gI, v);
                        uint4 o;
                        o.x = pack_h2(v[0], v[1]); o.y = pack_h2(v[2], v[3]); o.z = pack_h2(v[4], v[5]); o.w = pack_h2(v[6], v[7]);
                        if (!(a.dbg & 4)) *reinterpret_cast<uint4 *>(Ct + (size_t)((col >> 3) + gI) * (kTileM * 16)) = o;
                    }
                }
            } else {
                // thread = row: park the 32 columns in shared memory, then every lane writes one column of each
                // row so a row's 32 floats leave as one 128-byte store
#pragma unroll
                for (int gI = 0; gI < 4; ++gI) {
                    if (gI < ng) {
                        float v[8];
                        act8(r + 8 * gI, col + 8 * gI, v);
#pragma unroll
                        for (int i = 0; i < 8; ++i) tr[lane * 33 + 8 * gI + i] = v[i];
                    }
                }
                __syncwarp();
                const int cl = col + lane;
                if (cl < a.n_valid && lane < 8 * ng) {
                    float *gp = g0 + cl;
                    for (int rr = 0; rr < rows_valid; ++rr) gp[(size_t)rr * (kFrames * kBins)] = tr[rr * 33 + lane];
                }
                __syncwarp();
            }
        };
        auto load = [&](uint32_t taddr, int c, int NT, uint32_t (&r)[32]) {
            if (c + 32 <= NT) {
                tmem_ld32(taddr + c, r);
            } else {  // N-tiles are multiples of 16
                uint32_t r16[16];
                tmem_ld16(taddr + c, r16);
#pragma unroll
                for (int i = 0; i < 16; ++i) { r[i] = r16[i]; r[16 + i] = 0; }
            }
        };

        for (int mp = pair0; mp < n_mpairs; mp += pair_stride, ++im) {
            const int mt = 2 * mp + (int)rank;
            const bool valid = mt < a.mt_count;
            const int q = mt / a.T, t = mt % a.T;
            if (OUT == 0) {
                Ct = reinterpret_cast<uint8_t *>(a.C) + (size_t)(q * a.C_q + t + a.C_t0) * c_tile_bytes +
                     (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16;
            } else {
                // gains are [sc][50*Kchunks][161] and sequence g = sc*Kchunks + jc owns frames 50*jc.. of stream-channel
                // sc, so its frame t is row 50*g + t of the flat [.][161] array: consecutive sequences are 50 rows apart
                const int gq = a.seq0 + q * kTileM + wq * 32;
                rows_valid = min(32, a.n_seq - gq);
                g0 = a.gains + ((size_t)kFrames * gq + (t + a.C_t0)) * kBins;
            }
            for (int j = 0; j < a.n_tiles; ++j, ++work) {
                const int NT = a.nt_size[j], col0 = a.nt_col[j];
                const uint32_t buf = work % n_acc;
                mbar_wait(&sh->acc_full[buf], (work / n_acc) & 1);
                tc_fence_after();
                if (warp == 4 && lane == 0) PAIR_TRACE(3, im, j, 0);
                const uint32_t tbase = tmem + buf * (uint32_t)pl.acc_cols + ((uint32_t)(wq * 32) << 16);
                // The loads run one chunk ahead of the arithmetic (tcgen05.wait::ld covers every load issued so far, so the
                // next chunk's load is issued right after the wait for the current one), and the accumulator goes back to
                // the MMA issuer as soon as this warp's last chunk sits in registers.
                const int NTv = (valid && !(a.dbg & 1)) ? NT : 0;  // (dbg 1, experiments: the epilogue drains nothing)
                constexpr int kStride = 32 * NSETS;
                int c = hset * 32;
                bool released = false;
                auto release = [&]() {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_addr(acc_empty_leader + 8u * buf);
                    released = true;
                };
                if (NSETS == 1) {
                    uint32_t rA[32], rB[32];
                    if (c < NTv) load(tbase, c, NT, rA);
                    while (c < NTv) {
                        tmem_wait_ld();
                        int c2 = c + kStride;
                        if (c2 < NTv) load(tbase, c2, NT, rB); else release();
                        emit(rA, col0 + c, c + 32 <= NT ? 4 : 2);
                        c = c2;
                        if (c >= NTv) break;
                        tmem_wait_ld();
                        c2 = c + kStride;
                        if (c2 < NTv) load(tbase, c2, NT, rA); else release();
                        emit(rB, col0 + c, c + 32 <= NT ? 4 : 2);
                        c = c2;
                    }
                } else {  // two warp sets overlap each other's loads and arithmetic: one register buffer per warp is enough
                    uint32_t r[32];
                    for (; c < NTv; c += kStride) {
                        load(tbase, c, NT, r);
                        tmem_wait_ld();
                        if (c + kStride >= NTv) release();
                        emit(r, col0 + c, c + 32 <= NT ? 4 : 2);
                    }
                }
                if (!released) release();
                if (warp == 4 && lane == 0) PAIR_TRACE(3, im, j, 1);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();  // neither CTA frees TMEM or exits while the other may still signal it
    if (warp == 2) tmem_dealloc2<512>(tmem);
}

// cuTensorMapEncodeTiled through the runtime (the library does not link libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        FVAD_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
        if (qres != cudaDriverEntryPointSuccess || !p) throw Error(FVAD_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
        fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

}  // namespace

// cuTensorMapEncodeTiled with the settings every map of this backend uses (no interleave, no swizzle, zero fill)
void encode_tiled_map(CUtensorMap *m, CUtensorMapDataType dt, unsigned rank, void *base, const cuuint64_t *gdim, const cuuint64_t *gstr,
                      const cuuint32_t *box) {
    const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    const CUresult rc = encode_tiled()(m, dt, rank, base, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) throw Error(FVAD_ERR_CUDA, "cuTensorMapEncodeTiled failed (" + std::to_string((int)rc) + ")");
}

// K-steps per weight slab of the pair kernel: equal pieces of at most 8 K-steps (the issue loop's unroll)
int pair_slab_ksteps(int K) {
    const int nk = K / 16;
    const int n_slabs = (nk + 7) / 8;
    return (nk + n_slabs - 1) / n_slabs;
}

// Tensor maps of a layer's CTA-pair weight image: N-tile j, half r = a [K/8 core columns][NT/2 rows x 16 B] array at
// img2 + nt_off[j] + r * (NT/2)*K*2; one box = the (NT/2 rows) x (ks K-steps) piece a CTA stores of a weight stage.
void build_pair_maps(PairMaps &maps, const void *img2, int K, int n_tiles, const int *nt_size, const unsigned *nt_off) {
    const int ks = pair_slab_ksteps(K);
    for (int j = 0; j < n_tiles; ++j)
        for (int r = 0; r < 2; ++r) {
            const int NH = nt_size[j] / 2;
            if (nt_size[j] % 16 || NH % 8 || NH * 2 > 256) throw Error(FVAD_ERR_INVALID_ARGUMENT, "pair GEMM: N-tile must be a multiple of 16, at most 256");
            void *base = const_cast<uint8_t *>(static_cast<const uint8_t *>(img2)) + nt_off[j] + (size_t)r * NH * K * 2;
            const cuuint64_t gdim[2] = {(cuuint64_t)NH * 2, (cuuint64_t)K / 8};  // 8-byte elements: a core column is NH*16 bytes
            const cuuint64_t gstr[1] = {(cuuint64_t)NH * 16};
            const cuuint32_t box[2] = {(cuuint32_t)NH * 2, (cuuint32_t)(2 * ks)};
            encode_tiled_map(&maps.m[2 * j + r], CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, base, gdim, gstr, box);
        }
}

template <int ACT, int OUT, int NSETS, bool GATHER = false>
static void launch_pair_variant(const GemmArgs &a, const PairPlan &pl, const PairMaps &maps, int grid, cudaStream_t s) {
    static bool attr = false;
    if (!attr) {
        FVAD_CUDA_CHECK(cudaFuncSetAttribute(tc_gemm_pair_kernel<ACT, OUT, NSETS, GATHER>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
        attr = true;
    }
    tc_gemm_pair_kernel<ACT, OUT, NSETS, GATHER><<<grid, 384 + 128 * (NSETS - 1), pl.smem_bytes, s>>>(a, pl, maps);
}

void launch_tc_gemm2(const GemmArgs &a, const PairMaps &maps, int n_sm, cudaStream_t s) {
    if (a.mt_count <= 0) return;
    PairPlan pl;
    int nt_max = 0;
    for (int j = 0; j < a.n_tiles; ++j) nt_max = nt_max > a.nt_size[j] ? nt_max : a.nt_size[j];
    if (a.K % 16 || a.K / 2 + 2 * nt_max > 512)
        throw Error(FVAD_ERR_INVALID_ARGUMENT, "pair GEMM: A tile + two accumulators exceed the 512 TMEM columns");
    // A third accumulator when the N-tiles leave room for it (not with the N-tiles nn_tc.cu cuts today: measured
    // per 20 stream-hours, W at N-tiles of 96 x three accumulators 7.4 ms against 144 x two 7.1).
    pl.n_acc = a.K / 2 + 3 * nt_max <= 512 ? 3 : 2;
    pl.nk_total = a.K / 16;
    pl.ks = pair_slab_ksteps(a.K);
    pl.n_slabs = (pl.nk_total + pl.ks - 1) / pl.ks;
    pl.acc_cols = nt_max;
    pl.stage_bytes = (nt_max / 2) * 16 * 2 * pl.ks;
    pl.a_slab_bytes = pl.ks * 2 * kTileM * 16;
    const int tr_bytes = a.out_mode == 1 ? 8 * 32 * 33 * 4 : 0;
    static_assert(sizeof(PairShared) <= 1024, "barrier block");
    const int fixed = ((a.Nc * 4 + 15) / 16) * 16 + tr_bytes + 1024 + 16;
    // A staging: a whole tile's slabs when that leaves the weight ring six stages.  Timeline: with a buffer fewer the last slab's load starts only when the first
    // buffer has been restaged and lands ~1 000 cycles late; with a ring that holds less than two N-tiles the producer
    // waits for MMA completions and the L2 latency is exposed.
    int min_stages = 8;  // (A/B per 20 stream-hours: fc3 5.22 -> 4.84 ms with 9 stages + 3 A buffers instead of 6 + 4; K = 400 layers indifferent)
    if (const char *e = getenv("FVAD_PAIR_MINSTAGES")) min_stages = atoi(e);  // experiments
    if (a.gather_rows && (a.K != 16 * kGatherKsteps || pl.ks != kGatherKs || pl.n_slabs != 2 || a.out_mode != 0 || a.act != 0))
        throw Error(FVAD_ERR_INVALID_ARGUMENT, "pair GEMM: the row-gather form is built for the K = 176 projection only");
    pl.n_abuf = a.gather_rows ? 0 : pl.n_slabs < kMaxABuf ? pl.n_slabs : kMaxABuf;
    while (pl.n_abuf > 2 && (kSmemTotal - pl.n_abuf * pl.a_slab_bytes - fixed) / pl.stage_bytes < min_stages) --pl.n_abuf;
    pl.b_off = pl.n_abuf * pl.a_slab_bytes;
    pl.n_stages = (kSmemTotal - pl.b_off - fixed) / pl.stage_bytes;
    if (pl.n_stages > kMaxStages) pl.n_stages = kMaxStages;
    if (pl.n_stages < pl.n_slabs + 1 || pl.n_slabs > kMaxSlabs) throw Error(FVAD_ERR_INVALID_ARGUMENT, "pair GEMM: layer does not fit the shared-memory plan");
    pl.bias_off = pl.b_off + pl.n_stages * pl.stage_bytes;
    pl.tr_off = pl.bias_off + ((a.Nc * 4 + 15) / 16) * 16;
    pl.sh_off = (pl.tr_off + tr_bytes + 15) & ~15;
    pl.smem_bytes = pl.sh_off + 1024;
    const int n_mpairs = (a.mt_count + 1) / 2;
    const int grid = 2 * (n_mpairs < n_sm / 2 ? n_mpairs : n_sm / 2);
    if (a.out_mode == 1) {
        if (a.act != 2) throw Error(FVAD_ERR_INVALID_ARGUMENT, "tc_gemm: the gains output applies the sigmoid");
        launch_pair_variant<2, 1, 2>(a, pl, maps, grid, s);
    } else if (a.act == 0 && a.gather_rows) launch_pair_variant<0, 0, 2, true>(a, pl, maps, grid, s);
    else if (a.act == 0) launch_pair_variant<0, 0, 2>(a, pl, maps, grid, s);
    else if (a.act == 1) launch_pair_variant<1, 0, 2>(a, pl, maps, grid, s);
    else throw Error(FVAD_ERR_INVALID_ARGUMENT, "tc_gemm: unsupported activation/output combination");
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace tc
}  // namespace fvad
