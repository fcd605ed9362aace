// nn_tc_gru3.cu -- the FIRST GRU layer with its input projection fused in (CTA pairs, tcgen05 cta_group::2).
//
// Replaces, inside `onnx_instance.run()` (reference src/NSNet2.zig:220): the fc1 MatMul/Add, the first GRU node's W x + Wb
// and its recurrence (ONNX GRU, linear_before_reset = 1, gate order z,r,h -- SURVEY App. C).  Used when the graph has no
// activation between fc1 and the GRU, so that fc1 folds into W (nn_tc.cu): the layer's input is then the 161 log-power
// features themselves (K = 176 with padding), and per step and per chunk of 16 hidden units
//     acc[nh | z | r | nx]  =  [0 | Wz' x | Wr' x | Wn' x]  +  [Rn h | Rz h | Rr h | 0]
// is ONE accumulator of 64 columns filled by 11 + 25 MMAs: the x part (B = the folded W rows [z r n], N = 48, columns
// 16..63, first MMA overwrites) and the h part (B = R rows [n z r], N = 48, columns 0..47, every MMA accumulates; columns
// 0..15 are zeroed by the epilogue when it drains the accumulator).  The n gate needs Wn x and Rn h apart
// (n = tanh(Wn x + b + r * (Rn h + Rbn))), which is what the two outer column groups are for.
//
// Why (profiles/r02_gru_knockouts.txt): the two-kernel form writes the projection W x (1200 fp16 per step-row, 43 GB per
// 20 x 2 h step) to HBM and the recurrent kernel's epilogue reads it back -- 12 loads and 48 fp16->fp32 converts per thread
// per chunk on the very path (MIO) that slows the tensor pipe's operand fetch -- while the tensor pipe itself has slack.
// Here the projection costs 11 more MMAs per chunk and nothing else: no projection kernel, no round trip, no pack kernel
// (the step's feature rows go from the STFT's fp16 rows straight into TMEM), and the x part of a step's first chunks is
// issued while the previous step's last hidden units are still being computed (it does not depend on h).
//
// Biases ride in the padding: input columns 161 and 162 are set to 1.0 and the folded W carries the bias there as an fp16
// hi + lo pair (exact to 2^-22 relative).  z and r rows are stored halved (sigma(x) = 0.5 + 0.5 tanh(x/2)).
//
// Roles (512 threads per CTA; both CTAs run all but the issuers): warp 0 weight producer (TMA: R half + W half per stage,
// bytes complete on the leader's barrier), warps 1 and 3 MMA issuers (leader; even / odd chunks, one accumulator each),
// warp 2 TMEM allocator, warps 4..7 / 8..11 epilogue of the even / odd chunks, warps 12..15 x stagers (thread = row: the
// next step's feature row waits in registers and goes into the free x buffer with tcgen05.st).
// TMEM columns: h [0,200) | x ping [200,288) | x pong [288,376) | acc0 [376,440) | acc1 [440,504).
// h is single-buffered: h_t is parked in shared memory (fp16) until the step's last MMAs have retired (nn_tc_gru2.cu).
#include "nn_tc.cuh"
#include "tc_dev.cuh"

namespace fvad {
namespace tc {

namespace {

constexpr int kH = 400;
constexpr int kKx = 176;                            // padded input width (161 features + 2 bias columns + zeros)
constexpr int kXSteps = kKx / 16;                   // 11
constexpr int kChunkUnits = 16;
constexpr int kChunkRows = 48;                      // N of both MMA forms
constexpr int kHalfRows = kChunkRows / 2;           // B rows held per CTA
constexpr int kChunks = kH / kChunkUnits;           // 25
constexpr int kRHalfBytes = kHalfRows * kH * 2;     // 19 200
constexpr int kWHalfBytes = kHalfRows * kKx * 2;    //  8 448
constexpr int kStageBytes = kRHalfBytes + kWHalfBytes;
constexpr int kStages = 4;
constexpr int kThreads = 512;
constexpr int kHsBytes = kChunks * kTileM * 32;     // h_t as fp16, [chunk][row][16 units]
constexpr int kSmem = kStages * kStageBytes + kHsBytes + kH * 4 + 1024;
constexpr uint32_t kColH = 0, kColX = 200, kColAcc = 376, kAccCols = 64;

struct Gru3Shared {
    uint64_t full[kStages];   // LEADER's: both CTAs' halves of a stage (R + W) complete their bytes here
    uint64_t empty[kStages];  // per CTA (multicast commit)
    uint64_t acc_full[2];     // per CTA (multicast commit)
    uint64_t acc_empty[2];    // LEADER's: 4 epilogue warps of the group in each CTA; first arrival = the initial zero fill
    uint64_t step_done;       // per CTA: every MMA of the step has retired (one commit per issuer)
    uint64_t x_free[2];       // per CTA: the MMAs that read x buffer b have retired (one commit per issuer, every second step)
    uint64_t h_ready;         // LEADER's: h_t is in both CTAs' TMEM (8 + 8 epilogue warps)
    uint64_t x_ready[2];      // LEADER's: x_t is in both CTAs' TMEM buffer b (4 + 4 stager warps)
    uint32_t tmem_base;
    uint32_t passed[2];       // LEADER's: per issuer, 1 + the last chunk (producer's sequence) whose full-wait it has passed
};

__device__ __forceinline__ float tanh_hw(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float2 tanh2(float2 x) { return make_float2(tanh_hw(x.x), tanh_hw(x.y)); }

__device__ __forceinline__ void tmem_st16_zero(uint32_t taddr) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1,%1};" ::"r"(taddr), "r"(0u)
        : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
    tc_gru3_kernel(const Gru3Args a, int n_pairs_total, const __grid_constant__ CUtensorMap rmap, const __grid_constant__ CUtensorMap wmap) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t *hs = smem + kStages * kStageBytes;
    float *rbn_s = reinterpret_cast<float *>(hs + kHsBytes);
    Gru3Shared *sh = reinterpret_cast<Gru3Shared *>(hs + kHsBytes + kH * 4);
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
            mbar_init(&sh->acc_empty[i], 8);
            mbar_init(&sh->x_free[i], 2);
            mbar_init(&sh->x_ready[i], 8);
        }
        mbar_init(&sh->step_done, 2);
        mbar_init(&sh->h_ready, 16);
        sh->passed[0] = sh->passed[1] = 0;
        fence_mbar_init();
    }
    for (int i = threadIdx.x; i < kH; i += kThreads) rbn_s[i] = a.rbn[i];
    if (warp == 2) tmem_alloc2<512>(&sh->tmem_base);
    tc_fence_before();
    __syncthreads();
    cluster_sync();  // both CTAs' barriers and TMEM exist before any remote arrive, remote complete_tx or paired MMA
    tc_fence_after();
    const uint32_t tmem = sh->tmem_base;

    const int n_p = (n_pairs_total - pair0 + n_pair_ctas - 1) / n_pair_ctas;  // tile pairs this cluster processes
    const uint32_t n_fill = (uint32_t)n_p * kSteps * kChunks;

    if (warp == 0) {
        // ===== producer: this CTA's halves (24 rows of R, 24 rows of the folded W) of every chunk =====
        if (elect_one()) {
            const uint32_t full_leader = mapa_u32(&sh->full[0], 0);
            uint32_t it = 0, st = 0, ph = 0;
            for (int ip = 0; ip < n_p; ++ip)
                for (int t = 0; t < kSteps; ++t)
                    for (int c = 0; c < kChunks; ++c, ++it) {
                        mbar_wait(&sh->empty[st], ph ^ 1);
                        if (leader) mbar_expect_tx(&sh->full[st], 2 * kStageBytes);
                        tma2_load_3d(smem + st * kStageBytes, &rmap, full_leader + 8u * st, 0, 0, 2 * c + (int)rank);
                        tma2_load_3d(smem + st * kStageBytes + kRHalfBytes, &wmap, full_leader + 8u * st, 0, 0, 2 * c + (int)rank);
                        if (++st == (uint32_t)kStages) { st = 0; ph ^= 1; }
                    }
            // drain: the leader's multicast commits to this CTA's `empty` barriers must have landed before exit
            for (uint32_t s = 0; s < (uint32_t)kStages && s < n_fill; ++s) {
                const uint32_t last = ((n_fill - 1 - s) / kStages) * kStages + s;
                mbar_wait(&sh->empty[s], (last / kStages) & 1);
            }
        }
    } else if (warp == 1 || warp == 3) {
        // ===== MMA issuers (leader only): warp 1 the even chunks (accumulator 0), warp 3 the odd ones =====
        if (leader && elect_one()) {
            constexpr uint32_t idesc = make_idesc_f16(2 * kTileM, kChunkRows);
            const uint32_t me = warp == 1 ? 0u : 1u;
            const uint32_t d = tmem + kColAcc + me * kAccCols;
            const uint32_t hT = tmem + kColH;
            volatile uint32_t *passed_me = &sh->passed[me], *passed_other = &sh->passed[me ^ 1u];
            uint32_t step = 0, use = 0;
            for (int ip = 0; ip < n_p; ++ip)
                for (int t = 0; t < kSteps; ++t, ++step) {
                    const uint32_t xb = step & 1u;
                    const uint32_t xT = tmem + kColX + xb * (kKx / 2);
                    for (int c = (int)me; c < kChunks; c += 2) {
                        const uint32_t it = step * kChunks + (uint32_t)c;
                        const uint32_t st = it % kStages, ph = (it / kStages) & 1;
                        // (a parity wait tells "this fill" from "the fill before" only: see nn_tc_gru2.cu)
                        if (it >= (uint32_t)kStages && (((it - kStages) % kChunks) & 1u) != me) {
                            while (*passed_other <= it - kStages) {}
                            __threadfence_block();
                        }
                        mbar_wait(&sh->full[st], ph);
                        __threadfence_block();
                        *passed_me = it + 1u;
                        const uint64_t rd0 = make_smem_desc(smem_u32(smem + st * kStageBytes), kHalfRows * 16, 128);
                        const uint64_t wd0 = make_smem_desc(smem_u32(smem + st * kStageBytes + kRHalfBytes), kHalfRows * 16, 128);
                        const uint32_t rlo = (uint32_t)rd0, rhi = (uint32_t)(rd0 >> 32), wlo = (uint32_t)wd0, whi = (uint32_t)(wd0 >> 32);
                        mbar_wait(&sh->acc_empty[me], use & 1);  // (the first use waits for the epilogue's initial zero fill)
                        ++use;
                        if (c < 2) mbar_wait(&sh->x_ready[xb], (step >> 1) & 1);  // x_t is in both CTAs' TMEM
                        tc_fence_after();
                        // x part: columns 16..63 <- [Wz' | Wr' | Wn'] x_t, the first MMA overwrites
                        mma2_ts_lohi<false>(d + 16, xT, wlo, whi, idesc);
#pragma unroll
                        for (int kk = 1; kk < kXSteps; ++kk) mma2_ts_lohi<true>(d + 16, xT + kk * 8, wlo + kk * (2 * kHalfRows), whi, idesc);
                        if (c < 2) {  // h_{t-1} (or the zero fill) is in both CTAs' TMEM
                            mbar_wait(&sh->h_ready, step & 1);
                            tc_fence_after();
                        }
                        // h part: columns 0..47 += [Rn | Rz | Rr] h_{t-1} (columns 0..15 were zeroed by the epilogue)
#pragma unroll
                        for (int kk = 0; kk < kH / 16; ++kk) mma2_ts_lohi<true>(d, hT + kk * 8, rlo + kk * (2 * kHalfRows), rhi, idesc);
                        mma2_commit_mc(&sh->empty[st], 3);
                        mma2_commit_mc(&sh->acc_full[me], 3);
                    }
                    mma2_commit_mc(&sh->step_done, 3);  // every MMA of this thread that reads h_{t-1} / x_t has retired when this lands
                    mma2_commit_mc(&sh->x_free[xb], 3);
                }
        }
    } else if (warp >= 12) {
        // ===== x stagers: thread = row; the step's feature row -> the free x buffer =====
        const int row = (warp & 3) * 32 + lane;
        const uint32_t xrow = tmem + kColX + ((uint32_t)((warp & 3) * 32) << 16);
        const uint32_t x_ready_leader = mapa_u32(&sh->x_ready[0], 0);
        uint32_t step = 0;
        for (int ip = 0; ip < n_p; ++ip) {
            const int q = 2 * (pair0 + ip * n_pair_ctas) + (int)rank;
            const int g = a.seq0 + q * kTileM + row;
            const bool ok = q < a.Q && g < a.n_seq;
            const int sc = ok ? g / a.Kchunks : 0, j = ok ? g - sc * a.Kchunks : 0;
            // the sequence's 54-row window: rows 50 j .. 50 j + 53 of its stream-channel's fp16 feature rows
            const uint4 *src = reinterpret_cast<const uint4 *>(a.feat16 + ((size_t)sc * (kCtx + kFrames * a.Kchunks) + (size_t)kFrames * j) * kKx);
            uint4 d[2 * kXSteps];
            auto fetch = [&](int t) {
#pragma unroll
                for (int i = 0; i < 2 * kXSteps; ++i) d[i] = ok ? __ldg(src + (size_t)t * (kKx / 8) + i) : make_uint4(0u, 0u, 0u, 0u);
                // bias columns 161 and 162 = 1.0 (halves 161, 162 sit in d[20].x high and d[20].y low)
                d[20].x = (d[20].x & 0x0000FFFFu) | 0x3C000000u;
                d[20].y = (d[20].y & 0xFFFF0000u) | 0x00003C00u;
            };
            fetch(0);
            for (int t = 0; t < kSteps; ++t, ++step) {
                const uint32_t xb = step & 1u;
                if (step >= 2) mbar_wait(&sh->x_free[xb], ((step >> 1) - 1) & 1);  // the MMAs of two steps ago have read this buffer
                tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < kXSteps; ++kk) {
                    const uint32_t r[8] = {d[2 * kk].x, d[2 * kk].y, d[2 * kk].z, d[2 * kk].w, d[2 * kk + 1].x, d[2 * kk + 1].y, d[2 * kk + 1].z, d[2 * kk + 1].w};
                    tmem_st8(xrow + xb * (kKx / 2) + kk * 8, r);
                }
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_addr(x_ready_leader + 8u * xb);
                if (t + 1 < kSteps) fetch(t + 1);
            }
        }
        // drain the leader's last multicast commits to this CTA's x_free barriers
        for (uint32_t b = 0; b < 2; ++b) {
            if (step > b) {
                const uint32_t last = ((step - 1 - b) / 2);  // completions of x_free[b] so far, minus one
                mbar_wait(&sh->x_free[b], last & 1);
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue: group g = even / odd chunks; rows of THIS CTA's tile =====
        const int g = (warp - 4) >> 2;
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        const uint32_t lane_off = (uint32_t)(wq * 32) << 16;
        const size_t row_off = (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16;
        const uint32_t acc_empty_leader = mapa_u32(&sh->acc_empty[g], 0);
        const uint32_t h_ready_leader = mapa_u32(&sh->h_ready, 0);
        const uint32_t th = tmem + lane_off + kColH;
        const uint32_t acc = tmem + lane_off + kColAcc + (uint32_t)g * kAccCols;
        uint4 *hrow = reinterpret_cast<uint4 *>(hs + (size_t)row * 32);  // + chunk * (kTileM * 32) bytes
        uint32_t step = 0, use = 0;
        // the accumulator's nh columns start at zero (the h part only ever accumulates)
        tmem_st16_zero(acc);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_addr(acc_empty_leader);
        for (int ip = 0; ip < n_p; ++ip) {
            const int q = 2 * (pair0 + ip * n_pair_ctas) + (int)rank;
            const bool valid = q < a.Q;  // an odd tile count leaves the last pair's second tile as padding
            {   // h_{-1} = 0 (the previous tile's last MMAs have retired: the step_done wait below)
                uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                for (int c = g; c < kChunks; c += 2) tmem_st8(th + c * 8, z);
                tmem_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_addr(h_ready_leader);
            }
            for (int t = 0; t < kSteps; ++t, ++step) {
                uint8_t *hot = reinterpret_cast<uint8_t *>(a.hout) + (size_t)(q * kSteps + t) * tile_bytes(kH) + row_off;
                const bool store = valid && t >= a.t_store_min;
                for (int c = g; c < kChunks; c += 2) {
                    mbar_wait(&sh->acc_full[g], use & 1);
                    ++use;
                    tc_fence_after();
                    uint32_t anh[16], az[16], ar[16], anx[16], ho[8];
                    tmem_ld16(acc, anh);
                    tmem_ld16(acc + 16, az);
                    tmem_ld16(acc + 32, ar);
                    tmem_ld16(acc + 48, anx);
                    tmem_ld8(th + c * 8, ho);  // h_{t-1} of this row's 16 units: TMEM keeps it untouched for the whole step
                    tmem_wait_ld();
                    tmem_st16_zero(acc);       // the next chunk's h part accumulates onto zeros
                    tmem_wait_st();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_addr(acc_empty_leader);  // drained: this group's next chunk may overwrite it

                    float rb[16];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 v = *reinterpret_cast<const float4 *>(rbn_s + c * kChunkUnits + 4 * i);
                        rb[4 * i] = v.x; rb[4 * i + 1] = v.y; rb[4 * i + 2] = v.z; rb[4 * i + 3] = v.w;
                    }
                    const float2 kHalf = make_float2(0.5f, 0.5f), kNeg = make_float2(-1.0f, -1.0f);
                    uint32_t hn[8];
#pragma unroll
                    for (int p = 0; p < 8; ++p) {
                        const float2 hp = unpack_h2(ho[p]);
                        const float2 z = __ffma2_rn(tanh2(make_float2(__uint_as_float(az[2 * p]), __uint_as_float(az[2 * p + 1]))), kHalf, kHalf);
                        const float2 r = __ffma2_rn(tanh2(make_float2(__uint_as_float(ar[2 * p]), __uint_as_float(ar[2 * p + 1]))), kHalf, kHalf);
                        const float2 qn = __fadd2_rn(make_float2(__uint_as_float(anh[2 * p]), __uint_as_float(anh[2 * p + 1])), make_float2(rb[2 * p], rb[2 * p + 1]));
                        const float2 n = tanh2(__ffma2_rn(r, qn, make_float2(__uint_as_float(anx[2 * p]), __uint_as_float(anx[2 * p + 1]))));
                        const float2 h = __ffma2_rn(z, __ffma2_rn(n, kNeg, hp), n);  // n + z (h_{t-1} - n)
                        hn[p] = pack_h2(h.x, h.y);
                    }
                    hrow[c * (kTileM * 2)] = make_uint4(hn[0], hn[1], hn[2], hn[3]);
                    hrow[c * (kTileM * 2) + 1] = make_uint4(hn[4], hn[5], hn[6], hn[7]);
                    if (store) {
                        uint8_t *p = hot + (size_t)(c * 2) * (kTileM * 16);
                        *reinterpret_cast<uint4 *>(p) = make_uint4(hn[0], hn[1], hn[2], hn[3]);
                        *reinterpret_cast<uint4 *>(p + kTileM * 16) = make_uint4(hn[4], hn[5], hn[6], hn[7]);
                    }
                }
                // the step's last MMAs (both issuers) have read h_{t-1}: its TMEM columns may now take h_t
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

// Tensor maps of the fused layer's weight images: per chunk [half 0 | half 1], each 24 rows tiled with R = 24;
// R image rows in [n z r] order over K = 400, folded-W image rows in [z r n] order over K = 176.
void build_gru3_maps(CUtensorMap &rmap, CUtensorMap &wmap, const void *Rimg3, const void *Wimg3) {
    {
        const cuuint64_t gdim[3] = {(cuuint64_t)kHalfRows * 2, (cuuint64_t)kH / 8, (cuuint64_t)2 * kChunks};  // 8-byte elements
        const cuuint64_t gstr[2] = {(cuuint64_t)kHalfRows * 16, (cuuint64_t)kRHalfBytes};
        const cuuint32_t box[3] = {(cuuint32_t)kHalfRows * 2, (cuuint32_t)kH / 8, 1};
        encode_tiled_map(&rmap, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void *>(Rimg3), gdim, gstr, box);
    }
    {
        const cuuint64_t gdim[3] = {(cuuint64_t)kHalfRows * 2, (cuuint64_t)kKx / 8, (cuuint64_t)2 * kChunks};
        const cuuint64_t gstr[2] = {(cuuint64_t)kHalfRows * 16, (cuuint64_t)kWHalfBytes};
        const cuuint32_t box[3] = {(cuuint32_t)kHalfRows * 2, (cuuint32_t)kKx / 8, 1};
        encode_tiled_map(&wmap, CU_TENSOR_MAP_DATA_TYPE_UINT64, 3, const_cast<void *>(Wimg3), gdim, gstr, box);
    }
}

void launch_tc_gru3(const Gru3Args &a, const CUtensorMap &rmap, const CUtensorMap &wmap, int n_sm, cudaStream_t s) {
    static bool attr = false;
    if (!attr) {
        FVAD_CUDA_CHECK(cudaFuncSetAttribute(tc_gru3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem));
        attr = true;
    }
    if (a.Q <= 0) return;
    const int n_pairs = (a.Q + 1) / 2;
    const int grid = 2 * (n_pairs < n_sm / 2 ? n_pairs : n_sm / 2);
    tc_gru3_kernel<<<grid, kThreads, kSmem, s>>>(a, n_pairs, rmap, wmap);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace tc
}  // namespace fvad
