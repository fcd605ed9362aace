// nn_tc_gemm.cu -- persistent warp-specialised tcgen05 GEMM with a fused bias/activation epilogue:
// the dense layers of NSNet2 (fc1, the GRU input projections W x + b for all 54 steps at once,
// fc2, fc3, fc4) over M-tiles of 128 (sequence, step) rows.
//
// Replaces the MatMul/Add/Relu/Sigmoid nodes that onnxruntime executes inside
// `onnx_instance.run()` (reference src/NSNet2.zig:220).
//
// Roles (256 threads): warp 0 producer (cp.async.bulk of A and B K-slabs into a 4-stage ring),
// warp 1 MMA issuer (one elected lane, tcgen05.mma kind::f16, accumulators in TMEM, two
// accumulator buffers of 256 columns), warp 2 TMEM allocator, warps 4..7 epilogue (tcgen05.ld ->
// bias -> activation -> fp16 tiles for the next layer, or the f32 gains scatter).
#include "nn_tc.cuh"
#include "tc_dev.cuh"

namespace fvad {
namespace tc {

namespace {

constexpr int kStages = 4;
constexpr int kSlabK = 64;                          // K columns per pipeline stage
constexpr int kABytes = kTileM * kSlabK * 2;        // 16 KB
constexpr int kBBytes = 256 * kSlabK * 2;           // 32 KB (N-tile <= 256)
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kGemmThreads = 256;
constexpr int kGemmSmem = kStages * kStageBytes + 1024;

struct GemmShared {
    uint64_t full[kStages], empty[kStages], acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(kGemmThreads, 1) tc_gemm_kernel(const GemmArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    GemmShared *sh = reinterpret_cast<GemmShared *>(smem + kStages * kStageBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&sh->full[i], 1);
            mbar_init(&sh->empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sh->acc_full[i], 1);
            mbar_init(&sh->acc_empty[i], 4);
        }
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc<512>(&sh->tmem_base);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = sh->tmem_base;

    const int n_slabs = (a.K + kSlabK - 1) / kSlabK;
    const size_t a_tile_bytes = tile_bytes(a.K);

    if (warp == 0) {
        // ===== producer =====
        if (elect_one()) {
            uint32_t it = 0;
            for (int mt = blockIdx.x; mt < a.mt_count; mt += gridDim.x) {
                const int q = mt / a.T, t = mt % a.T;
                const uint8_t *At = reinterpret_cast<const uint8_t *>(a.A) + (size_t)(q * a.A_q + t + a.A_t0) * a_tile_bytes;
                for (int j = 0; j < a.n_tiles; ++j) {
                    const int NT = a.nt_size[j];
                    const uint8_t *Bt = reinterpret_cast<const uint8_t *>(a.B) + a.nt_off[j];
                    for (int sl = 0; sl < n_slabs; ++sl, ++it) {
                        const int st = it % kStages;
                        mbar_wait(&sh->empty[st], ((it / kStages) & 1) ^ 1);
                        const int k0 = sl * kSlabK, kw = min(kSlabK, a.K - k0);
                        const uint32_t ab = (uint32_t)(kw / 8) * (kTileM * 16), bb = (uint32_t)(kw / 8) * (NT * 16);
                        uint8_t *sa = smem + st * kStageBytes, *sb = sa + kABytes;
                        mbar_expect_tx(&sh->full[st], ab + bb);
                        bulk_g2s(sa, At + (size_t)(k0 / 8) * (kTileM * 16), ab, &sh->full[st]);
                        bulk_g2s(sb, Bt + (size_t)(k0 / 8) * (NT * 16), bb, &sh->full[st]);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (elect_one()) {
            uint32_t it = 0, work = 0;
            for (int mt = blockIdx.x; mt < a.mt_count; mt += gridDim.x) {
                for (int j = 0; j < a.n_tiles; ++j, ++work) {
                    const int NT = a.nt_size[j];
                    const uint32_t buf = work & 1;
                    mbar_wait(&sh->acc_empty[buf], ((work >> 1) & 1) ^ 1);
                    tc_fence_after();
                    const uint32_t idesc = make_idesc_f16(kTileM, NT);
                    const uint32_t d = tmem + buf * 256;
                    for (int sl = 0; sl < n_slabs; ++sl, ++it) {
                        const int st = it % kStages;
                        mbar_wait(&sh->full[st], (it / kStages) & 1);
                        tc_fence_after();
                        const int kw = min(kSlabK, a.K - sl * kSlabK);
                        const uint32_t sa = smem_u32(smem + st * kStageBytes), sb = sa + kABytes;
                        for (int kk = 0; kk < kw / 16; ++kk) {
                            const uint64_t ad = make_smem_desc(sa + kk * 2 * (kTileM * 16), kTileM * 16, 128);
                            const uint64_t bd = make_smem_desc(sb + kk * 2 * (NT * 16), NT * 16, 128);
                            mma_ss(d, ad, bd, idesc, (sl > 0 || kk > 0) ? 1u : 0u);
                        }
                        mma_commit(&sh->empty[st]);  // frees the stage when these MMAs have read it
                    }
                    mma_commit(&sh->acc_full[buf]);
                }
            }
        }
    } else if (warp >= 4) {
        // ===== epilogue =====
        const int wq = warp & 3;
        const int row = wq * 32 + lane;
        const size_t c_tile_bytes = tile_bytes(a.Nc);
        uint32_t work = 0;
        for (int mt = blockIdx.x; mt < a.mt_count; mt += gridDim.x) {
            const int q = mt / a.T, t = mt % a.T;
            uint8_t *Ct = nullptr;
            float *grow = nullptr;
            if (a.out_mode == 0) {
                Ct = reinterpret_cast<uint8_t *>(a.C) + (size_t)(q * a.C_q + t + a.C_t0) * c_tile_bytes +
                     (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16;
            } else {
                const int g = a.seq0 + q * kTileM + row;
                if (g < a.n_seq) {
                    const int sc = g / a.Kchunks, jc = g % a.Kchunks;
                    grow = a.gains + ((size_t)sc * (kFrames * a.Kchunks) + (size_t)kFrames * jc + (t + a.C_t0)) * kBins;
                }
            }
            for (int j = 0; j < a.n_tiles; ++j, ++work) {
                const int NT = a.nt_size[j], col0 = a.nt_col[j];
                const uint32_t buf = work & 1;
                mbar_wait(&sh->acc_full[buf], (work >> 1) & 1);
                tc_fence_after();
                const uint32_t tbase = tmem + buf * 256 + ((uint32_t)(wq * 32) << 16);
                for (int c = 0; c < NT; c += 16) {
                    uint32_t r[16];
                    tmem_ld16(tbase + c, r);
                    tmem_wait_ld();
                    float v[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        float x = __uint_as_float(r[i]) + __ldg(a.bias + col0 + c + i);
                        if (a.act == 1) x = fmaxf(x, 0.0f);
                        else if (a.act == 2) x = __fdividef(1.0f, 1.0f + __expf(-x));
                        v[i] = x;
                    }
                    if (a.out_mode == 0) {
                        uint4 lo, hi;
                        lo.x = pack_h2(v[0], v[1]); lo.y = pack_h2(v[2], v[3]); lo.z = pack_h2(v[4], v[5]); lo.w = pack_h2(v[6], v[7]);
                        hi.x = pack_h2(v[8], v[9]); hi.y = pack_h2(v[10], v[11]); hi.z = pack_h2(v[12], v[13]); hi.w = pack_h2(v[14], v[15]);
                        uint8_t *p = Ct + (size_t)((col0 + c) >> 3) * (kTileM * 16);
                        *reinterpret_cast<uint4 *>(p) = lo;
                        *reinterpret_cast<uint4 *>(p + kTileM * 16) = hi;
                    } else if (grow) {
#pragma unroll
                        for (int i = 0; i < 16; ++i)
                            if (col0 + c + i < a.n_valid) grow[col0 + c + i] = v[i];
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&sh->acc_empty[buf]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc<512>(tmem);
}

}  // namespace

void launch_tc_gemm(const GemmArgs &a, int n_sm, cudaStream_t s) {
    static bool attr = false;
    if (!attr) {
        FVAD_CUDA_CHECK(cudaFuncSetAttribute(tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem));
        attr = true;
    }
    if (a.mt_count <= 0) return;
    const int grid = a.mt_count < n_sm ? a.mt_count : n_sm;
    tc_gemm_kernel<<<grid, kGemmThreads, kGemmSmem, s>>>(a);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace tc
}  // namespace fvad
