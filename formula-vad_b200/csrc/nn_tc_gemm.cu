// nn_tc_gemm.cu -- persistent warp-specialised tcgen05 GEMM with a fused bias/activation epilogue:
// the dense layers of NSNet2 (fc1, the GRU input projections W x + b for all 54 steps at once,
// fc2, fc3, fc4) over M-tiles of 128 (sequence, step) rows.
//
// Replaces the MatMul/Add/Relu/Sigmoid nodes that onnxruntime executes inside
// `onnx_instance.run()` (reference src/NSNet2.zig:220).
//
// Roles (256 threads): warp 0 producer (cp.async.bulk: the M-tile's A slabs once, B K-slabs into a ring),
// warp 1 MMA issuer (one elected lane, tcgen05.mma kind::f16, accumulators in TMEM, two
// accumulator buffers of 256 columns), warp 2 TMEM allocator, warps 4..7 epilogue (tcgen05.ld ->
// bias -> activation -> fp16 tiles for the next layer, or the f32 gains scatter).
#include "nn_tc.cuh"
#include "tc_dev.cuh"

namespace fvad {
namespace tc {

namespace {

// Shared memory: [A tile, all K columns | B ring: n_stages x (256 rows x KS cols max) | bias | barriers].
// The A tile of an M-tile is fetched once (through the same producer queue, slab by slab, while the
// first N-tile runs) and stays resident for all N-tiles, so L2 traffic per M-tile is A + the whole
// weight matrix: 128 flop per byte, the same intensity as the GRU kernel's R stream.
constexpr int kMaxSlabs = 40;
constexpr int kMaxStages = 16;
constexpr int kGemmThreads = 384;  // 4 control warps, 4 epilogue warps, 4 A-stager warps (TS only)
// 214 KB, not the full 227: the VAD scan of the previous span runs concurrently on its own stream (one warp and
// ~11 KB of shared memory per stream); a persistent GEMM CTA that takes a whole SM's shared memory cannot start
// on the SMs the scan occupies, and with a static tile assignment the late CTAs stretch the kernel by ~60 %
// (measured: W1, which overlaps the scan, 12.3 ms vs W2 7.6 ms for the same GEMM).
constexpr int kSmemTotal = 214 * 1024;

// timeline probe (CTA 0, first 4 M-tiles): [kind][mtile][ntile][slab], kinds: 0 producer B issued, 1 MMA full ok,
// 2 MMA issued, 3 epilogue acc_full ok (slab 0) / done (slab 1), 4 MMA acc_empty ok (slab 0)
#define GEMM_TRACE(kind, im, j, sl) do { if (a.trace && blockIdx.x == 0 && (im) < 4 && (j) < 8) \
    a.trace[(((kind) * 4 + (im)) * 8 + (j)) * 32 + (sl)] = (unsigned long long)clock64(); } while (0)

struct GemmShared {
    uint64_t full[kMaxStages], empty[kMaxStages], a_full[kMaxSlabs], a_empty[kMaxSlabs], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

struct GemmPlan {  // host-computed shared-memory plan
    int KS, n_slabs, n_stages, stage_bytes, a_bytes, bias_off, tr_off, sh_off, smem_bytes;
    int acc_cols;  // TMEM columns between the two accumulators
};

// ACT: 0 none, 1 relu, 2 sigmoid.  OUT: 0 fp16 tiles for the next layer, 1 f32 gains rows (coalesced through a
// per-warp shared-memory transpose).
//
// TS = true (GemmArgs::a_tmem): the A tile goes to TENSOR MEMORY instead.  Its K slabs travel through the ring like
// weight slabs; four stager warps (thread = row) move each from its ring stage into TMEM with tcgen05.st, slab by
// slab as the previous M-tile's last N-tile releases the columns (a_empty[sl] / a_full[sl]), and every MMA takes A
// from TMEM.  (tcgen05.cp from the MMA thread does the same copy without the extra warps but takes ~250 cycles per
// 128x32-byte piece, 5-7 k cycles per M-tile with the tensor pipe idle -- measured, then replaced.)  That frees A's 100 KB for the weight ring -- the SS form's timeline
// (profiles/r01_gemm_timeline.txt) is bound by ring bytes in flight, three 30 KB stages against a 1 400-2 400
// cycle L2 latency -- and halves the shared-memory read traffic per MMA.  TMEM columns: [acc 0 | acc 1 | A (K/2)].
template <int ACT, int OUT, bool TS>
__global__ void __launch_bounds__(kGemmThreads, 1) tc_gemm_kernel(const GemmArgs a, const GemmPlan pl) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *sA = smem, *sB = smem + pl.a_bytes;
    float *bias_s = reinterpret_cast<float *>(smem + pl.bias_off);
    GemmShared *sh = reinterpret_cast<GemmShared *>(smem + pl.sh_off);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < pl.n_stages; ++i) {
            mbar_init(&sh->full[i], 1);
            mbar_init(&sh->empty[i], 1);
        }
        for (int i = 0; i < pl.n_slabs; ++i) {
            mbar_init(&sh->a_full[i], 1);
            mbar_init(&sh->a_empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sh->acc_full[i], 1);
            mbar_init(&sh->acc_empty[i], 4);
        }
        fence_mbar_init();
    }
    for (int i = threadIdx.x; i < a.Nc; i += kGemmThreads) bias_s[i] = a.bias[i];
    if (warp == 2) tmem_alloc<512>(&sh->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sh->tmem_base;
    const int KS = pl.KS, n_slabs = pl.n_slabs, n_stages = pl.n_stages;
    const size_t a_tile_bytes = tile_bytes(a.K);

    if (warp == 0) {
        // ===== producer =====
        if (elect_one()) {
            uint32_t it = 0, st = 0, ph = 0, im = 0;
            for (int mt = blockIdx.x; mt < a.mt_count; mt += gridDim.x, ++im) {
                const int q = mt / a.T, t = mt % a.T;
                const uint8_t *At = reinterpret_cast<const uint8_t *>(a.A) + (size_t)(q * a.A_q + t + a.A_t0) * a_tile_bytes;
                if (TS) {  // this M-tile's A slabs, one ring stage each
                    for (int sl = 0; sl < n_slabs; ++sl) {
                        const int k0 = sl * KS, kw = min(KS, a.K - k0);
                        mbar_wait(&sh->empty[st], ph ^ 1);
                        const uint32_t ab = (uint32_t)(kw / 8) * (kTileM * 16);
                        mbar_expect_tx(&sh->full[st], ab);
                        bulk_g2s(sB + (size_t)st * pl.stage_bytes, At + (size_t)(k0 / 8) * (kTileM * 16), ab, &sh->full[st]);
                        if (++st == (uint32_t)n_stages) { st = 0; ph ^= 1; }
                    }
                }
                for (int j = 0; j < a.n_tiles; ++j) {
                    const int NT = a.nt_size[j];
                    const uint8_t *Bt = reinterpret_cast<const uint8_t *>(a.B) + a.nt_off[j];
                    for (int sl = 0; sl < n_slabs; ++sl, ++it) {
                        const int k0 = sl * KS, kw = min(KS, a.K - k0);
                        if (!TS && j == 0) {  // this M-tile's A slab (slot freed by the previous M-tile's last N-tile)
                            mbar_wait(&sh->a_empty[sl], (im & 1) ^ 1);
                            const uint32_t ab = (uint32_t)(kw / 8) * (kTileM * 16);
                            mbar_expect_tx(&sh->a_full[sl], ab);
                            bulk_g2s(sA + (size_t)(k0 / 8) * (kTileM * 16), At + (size_t)(k0 / 8) * (kTileM * 16), ab, &sh->a_full[sl]);
                        }
                        mbar_wait(&sh->empty[st], ph ^ 1);
                        const uint32_t bb = (uint32_t)(kw / 8) * (NT * 16);
                        if (a.dbg & 2) { mbar_arrive(&sh->full[st]); }
                        else {
                            mbar_expect_tx(&sh->full[st], bb);
                            bulk_g2s(sB + (size_t)st * pl.stage_bytes, Bt + (size_t)(k0 / 8) * (NT * 16), bb, &sh->full[st]);
                        }
                        GEMM_TRACE(0, im, j, sl);
                        if (++st == (uint32_t)n_stages) { st = 0; ph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            uint32_t st = 0, ph = 0, work = 0, im = 0;
            // descriptors as (lo, hi): the K loop advances lo by immediates (tc_dev.cuh, mma_ss_lohi)
            const uint64_t ad0 = make_smem_desc(smem_u32(sA), kTileM * 16, 128);
            const uint32_t alo0 = (uint32_t)ad0, ahi = (uint32_t)(ad0 >> 32);
            const uint32_t tA = tmem + 2u * (uint32_t)pl.acc_cols;  // TS: A image, 8 columns per K-step of 16
            for (int mt = blockIdx.x; mt < a.mt_count; mt += gridDim.x, ++im) {
                if (TS) {  // the stager warps consume this M-tile's A stages
                    st += (uint32_t)n_slabs;
                    while (st >= (uint32_t)n_stages) { st -= (uint32_t)n_stages; ph ^= 1; }
                }
                for (int j = 0; j < a.n_tiles; ++j, ++work) {
                    const int NT = a.nt_size[j];
                    const uint32_t buf = work & 1;
                    mbar_wait(&sh->acc_empty[buf], ((work >> 1) & 1) ^ 1);
                    tc_fence_after();
                    GEMM_TRACE(4, im, j, 0);
                    const uint32_t idesc = make_idesc_f16(kTileM, NT);
                    const uint32_t d = tmem + buf * (uint32_t)pl.acc_cols;
                    const bool last_n = (j == a.n_tiles - 1);
                    const uint32_t bstep = 2u * (uint32_t)NT;  // one K-step of B = 2 core columns = 2*NT*16 bytes
                    for (int sl = 0; sl < n_slabs; ++sl) {
                        if (j == 0) mbar_wait(&sh->a_full[sl], im & 1);
                        mbar_wait(&sh->full[st], ph);
                        tc_fence_after();
                        GEMM_TRACE(1, im, j, sl);
                        const int k0 = sl * KS, nk = min(KS, a.K - k0) / 16;
                        const uint64_t bd0 = make_smem_desc(smem_u32(sB + (size_t)st * pl.stage_bytes), NT * 16, 128);
                        const uint32_t blo = (uint32_t)bd0, bhi = (uint32_t)(bd0 >> 32);
                        const uint32_t alo = alo0 + (uint32_t)(k0 / 8) * (kTileM * 16 / 16);
                        if (TS) {
                            const uint32_t ta = tA + (uint32_t)(k0 / 16) * 8u;
                            if (sl == 0) mma_ts_lohi<false>(d, ta, blo, bhi, idesc);
                            else mma_ts_lohi<true>(d, ta, blo, bhi, idesc);
#pragma unroll
                            for (int kk = 1; kk < 8; ++kk)
                                if (kk < nk) mma_ts_lohi<true>(d, ta + kk * 8u, blo + kk * bstep, bhi, idesc);
                        } else {
                            if (sl == 0) mma_ss_lohi<false>(d, alo, ahi, blo, bhi, idesc);
                            else mma_ss_lohi<true>(d, alo, ahi, blo, bhi, idesc);
#pragma unroll
                            for (int kk = 1; kk < 4; ++kk)
                                if (kk < nk) mma_ss_lohi<true>(d, alo + kk * (2 * kTileM * 16 / 16), ahi, blo + kk * bstep, bhi, idesc);
                        }
                        mma_commit(&sh->empty[st]);               // B stage free once these MMAs have read it
                        if (last_n) mma_commit(&sh->a_empty[sl]); // A slab (smem, or its TMEM columns) free for the next M-tile
                        GEMM_TRACE(2, im, j, sl);
                        if (++st == (uint32_t)n_stages) { st = 0; ph ^= 1; }
                    }
                    mma_commit(&sh->acc_full[buf]);
                    GEMM_TRACE(4, im, j, 1);
                }
            }
        }
    } else if (warp >= 8) {
        // ===== A stagers (TS): ring stage -> TMEM, thread = row (lane quadrant = warp % 4) =====
        if (TS) {
            const int row = (warp & 3) * 32 + lane;
            const uint32_t tA = tmem + 2u * (uint32_t)pl.acc_cols + ((uint32_t)((warp & 3) * 32) << 16);
            uint32_t st = 0, ph = 0, im = 0;
            for (int mt = blockIdx.x; mt < a.mt_count; mt += gridDim.x, ++im) {
                for (int sl = 0; sl < n_slabs; ++sl) {
                    const int k0 = sl * KS, nk = min(KS, a.K - k0) / 16;
                    mbar_wait(&sh->a_empty[sl], (im & 1) ^ 1);  // the previous M-tile's MMAs over these columns are done
                    mbar_wait(&sh->full[st], ph);
                    tc_fence_after();
                    const uint8_t *src = sB + (size_t)st * pl.stage_bytes + (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16;
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        if (kk < nk) {
                            const uint4 lo = *reinterpret_cast<const uint4 *>(src + (size_t)(2 * kk) * (kTileM * 16));
                            const uint4 hi = *reinterpret_cast<const uint4 *>(src + (size_t)(2 * kk + 1) * (kTileM * 16));
                            const uint32_t r[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
                            tmem_st8(tA + (uint32_t)(k0 / 16 + kk) * 8u, r);
                        }
                    }
                    tmem_wait_st();
                    tc_fence_before();
                    bar_sync(1, 128);
                    if (warp == 8 && lane == 0) {
                        mbar_arrive(&sh->empty[st]);
                        mbar_arrive(&sh->a_full[sl]);
                    }
                    if (++st == (uint32_t)n_stages) { st = 0; ph ^= 1; }
                }
                st += (uint32_t)(a.n_tiles * n_slabs);
                while (st >= (uint32_t)n_stages) { st -= (uint32_t)n_stages; ph ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue =====
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        const size_t c_tile_bytes = tile_bytes(a.Nc);
        float *tr = reinterpret_cast<float *>(smem + pl.tr_off) + wq * (32 * 33);  // OUT == 1: this warp's 32x32 transpose pad
        uint32_t work = 0;
        int im = 0;

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

        for (int mt = blockIdx.x; mt < a.mt_count; mt += gridDim.x, ++im) {
            const int q = mt / a.T, t = mt % a.T;
            uint8_t *Ct = nullptr;
            float *g0 = nullptr;   // OUT == 1: gains row of this warp's first sequence
            int rows_valid = 0;
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
                const uint32_t buf = work & 1;
                mbar_wait(&sh->acc_full[buf], (work >> 1) & 1);
                tc_fence_after();
                if (warp == 4 && lane == 0) GEMM_TRACE(3, im, j, 0);
                const uint32_t tbase = tmem + buf * (uint32_t)pl.acc_cols + ((uint32_t)(wq * 32) << 16);
                for (int c = 0; c < ((a.dbg & 1) ? 0 : NT); c += 32) {
                    uint32_t r[32];
                    const bool full32 = c + 32 <= NT;  // N-tiles are multiples of 16
                    if (full32) {
                        tmem_ld32(tbase + c, r);
                    } else {
                        uint32_t r16[16];
                        tmem_ld16(tbase + c, r16);
#pragma unroll
                        for (int i = 0; i < 16; ++i) { r[i] = r16[i]; r[16 + i] = 0; }
                    }
                    tmem_wait_ld();
                    const int ng = full32 ? 4 : 2;  // groups of 8 columns
                    if (OUT == 0) {
#pragma unroll
                        for (int gI = 0; gI < 4; ++gI) {
                            if (gI < ng) {
                                float v[8];
                                act8(r + 8 * gI, col0 + c + 8 * gI, v);
                                uint4 o;
                                o.x = pack_h2(v[0], v[1]); o.y = pack_h2(v[2], v[3]); o.z = pack_h2(v[4], v[5]); o.w = pack_h2(v[6], v[7]);
                                *reinterpret_cast<uint4 *>(Ct + (size_t)(((col0 + c) >> 3) + gI) * (kTileM * 16)) = o;
                            }
                        }
                    } else {
                        // thread = row: park the 32 columns in shared memory, then every lane writes one column of each
                        // row so a row's 32 floats leave as one 128-byte store
#pragma unroll
                        for (int gI = 0; gI < 4; ++gI) {
                            if (gI < ng) {
                                float v[8];
                                act8(r + 8 * gI, col0 + c + 8 * gI, v);
#pragma unroll
                                for (int i = 0; i < 8; ++i) tr[lane * 33 + 8 * gI + i] = v[i];
                            }
                        }
                        __syncwarp();
                        const int col = col0 + c + lane;
                        if (col < a.n_valid && lane < 8 * ng) {
                            float *gp = g0 + col;
                            for (int rr = 0; rr < rows_valid; ++rr) gp[(size_t)rr * (kFrames * kBins)] = tr[rr * 33 + lane];
                        }
                        __syncwarp();
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (warp == 4 && lane == 0) GEMM_TRACE(3, im, j, 1);
                if (lane == 0) mbar_arrive(&sh->acc_empty[buf]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}

}  // namespace

template <int ACT, int OUT, bool TS>
static void launch_variant2(const GemmArgs &a, const GemmPlan &pl, int grid, cudaStream_t s) {
    static bool attr = false;
    if (!attr) {
        FVAD_CUDA_CHECK(cudaFuncSetAttribute(tc_gemm_kernel<ACT, OUT, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
        attr = true;
    }
    tc_gemm_kernel<ACT, OUT, TS><<<grid, kGemmThreads, pl.smem_bytes, s>>>(a, pl);
}
template <int ACT, int OUT>
static void launch_variant(const GemmArgs &a, const GemmPlan &pl, int grid, cudaStream_t s) {
    if (a.a_tmem) launch_variant2<ACT, OUT, true>(a, pl, grid, s);
    else launch_variant2<ACT, OUT, false>(a, pl, grid, s);
}

void launch_tc_gemm(const GemmArgs &a, int n_sm, cudaStream_t s) {
    if (a.mt_count <= 0) return;
    GemmPlan pl;
    int nt_max = 0;
    for (int j = 0; j < a.n_tiles; ++j) nt_max = nt_max > a.nt_size[j] ? nt_max : a.nt_size[j];
    pl.a_bytes = a.a_tmem ? 0 : a.K * 256;  // K/8 core columns x 2048 B
    pl.acc_cols = a.a_tmem ? nt_max : 256;
    if (a.a_tmem && (a.K / 2 + 2 * nt_max > 512 || a.K % 16))
        throw Error(FVAD_ERR_INVALID_ARGUMENT, "tc_gemm: A tile + two accumulators exceed the 512 TMEM columns");
    const int tr_bytes = a.out_mode == 1 ? 4 * 32 * 33 * 4 : 0;
    static_assert(sizeof(GemmShared) <= 1024, "barrier block");
    const int fixed = ((a.Nc * 4 + 15) / 16) * 16 + tr_bytes + 1024;
    pl.KS = a.a_tmem ? 128 : 64;  // (the SS issue loop is unrolled for 4 K-steps, the TS one for 8)
    if (a.a_tmem && (a.dbg & 8)) pl.KS = 64;
    for (;;) {
        pl.stage_bytes = (a.a_tmem && nt_max < kTileM ? kTileM : nt_max) * pl.KS * 2;  // (a stage also carries A slabs when a_tmem)
        pl.n_stages = (kSmemTotal - pl.a_bytes - fixed) / pl.stage_bytes;
        if (pl.n_stages >= 3 || pl.KS == 16) break;  // fewer, larger stages: the per-stage barrier round trip costs the issuing thread more than the shorter ring
        pl.KS /= 2;
    }
    if (pl.n_stages > kMaxStages) pl.n_stages = kMaxStages;
    pl.n_slabs = (a.K + pl.KS - 1) / pl.KS;
    if (pl.n_stages < 2 || pl.n_slabs > kMaxSlabs) throw Error(FVAD_ERR_INVALID_ARGUMENT, "tc_gemm: layer does not fit the shared-memory plan");
    // With A in TMEM two consumers share the ring and each skips the other's stages, while a parity wait can only tell
    // "this fill" from "the fill before".  The a_full / a_empty waits that precede every full-barrier wait guarantee the
    // previous fill of the awaited stage is over exactly when the ring is at least as long as an A tile's slab count.
    if (a.a_tmem && pl.n_stages < pl.n_slabs) throw Error(FVAD_ERR_INVALID_ARGUMENT, "tc_gemm: ring shorter than the A tile's slabs");
    pl.bias_off = pl.a_bytes + pl.n_stages * pl.stage_bytes;
    pl.tr_off = pl.bias_off + ((a.Nc * 4 + 15) / 16) * 16;
    pl.sh_off = pl.tr_off + tr_bytes;
    pl.smem_bytes = pl.sh_off + 1024;
    const int grid = a.mt_count < n_sm ? a.mt_count : n_sm;
    if (a.out_mode == 1) {
        if (a.act != 2) throw Error(FVAD_ERR_INVALID_ARGUMENT, "tc_gemm: the gains output applies the sigmoid");
        launch_variant<2, 1>(a, pl, grid, s);
    } else if (a.act == 0) launch_variant<0, 0>(a, pl, grid, s);
    else if (a.act == 1) launch_variant<1, 0>(a, pl, grid, s);
    else throw Error(FVAD_ERR_INVALID_ARGUMENT, "tc_gemm: unsupported activation/output combination");
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace tc
}  // namespace fvad
