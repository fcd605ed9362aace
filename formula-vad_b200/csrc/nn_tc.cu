// nn_tc.cu -- NSNet2 on the 5th-generation tensor cores: host side of the tcgen05 backend.
// Weight pre-tiling (fp16 core-matrix images, GRU gate rows re-ordered so that one 48-row chunk holds
// z|r|n of 16 units), the per-sub-batch schedule
//     pack -> fc1 -> W1 x -> GRU1 -> W2 h1 -> GRU2 -> fc2 -> fc3 -> fc4 (gains scatter)
// and the activation workspace.  Kernels: nn_tc_gemm.cu, nn_tc_gru.cu.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cstring>

#include "nn.cuh"
#include "nn_tc.cuh"

namespace fvad {

namespace tc {

// ---- pack kernel: features f32 -> fp16 tiles -------------------------------------------------
__global__ void __launch_bounds__(256) tc_pack_kernel(const PackArgs a) {
    // one CTA per (q, t) tile: 128 rows x 176 columns.  Row r of the tile is row 50j + t of its stream-channel's
    // fp16 feature rows (352 contiguous bytes): a warp fetches a row as 22 16-byte pieces and drops each into its
    // core column of a shared-memory image of the tile; the image then leaves with 16-byte stores in tile order.
    __shared__ __align__(16) uint8_t img[kTileM * 176 * 2];
    const int q = blockIdx.x / kSteps, t = blockIdx.x % kSteps;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int row = warp; row < kTileM; row += 8) {
        const int g = a.seq0 + q * kTileM + row;
        if (lane < 22) {
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (g < a.n_seq) {
                const int sc = g / a.Kchunks, j = g % a.Kchunks;
                v = __ldg(reinterpret_cast<const uint4 *>(a.feat16 + ((size_t)sc * (kCtx + kFrames * a.Kchunks) + (size_t)kFrames * j + t) * 176) + lane);
            }
            *reinterpret_cast<uint4 *>(img + (size_t)lane * (kTileM * 16) + (size_t)(row >> 3) * 128 + (size_t)(row & 7) * 16) = v;
        }
    }
    __syncthreads();
    uint4 *dst = reinterpret_cast<uint4 *>(reinterpret_cast<uint8_t *>(a.A0) + (size_t)blockIdx.x * tile_bytes(176));
    const uint4 *srcv = reinterpret_cast<const uint4 *>(img);
    for (int i = threadIdx.x; i < kTileM * 176 * 2 / 16; i += 256) dst[i] = srcv[i];
}

void launch_tc_pack(const PackArgs &a, cudaStream_t s) {
    if (a.Q <= 0) return;
    tc_pack_kernel<<<(unsigned)(a.Q * kSteps), 256, 0, s>>>(a);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace tc

namespace {

using tc::GemmArgs;

struct Layer {  // one dense layer prepared for launch_tc_gemm
    DevBuf<__half> img, img2;
    DevBuf<float> bias;
    int K = 0, Nc = 0, n_valid = 0, n_tiles = 0;
    int a_tmem = 0;  // GemmArgs::a_tmem for this layer (the N-tiles were cut for it)
    tc::PairMaps maps;  // CTA-pair kernel: tensor maps over img2 (built when the layer runs on pairs)
    bool pair = false;  // runs on CTA pairs (nn_tc_gemm2.cu)
    int nt_size[tc::kMaxNTiles] = {}, nt_col[tc::kMaxNTiles] = {};
    unsigned nt_off[tc::kMaxNTiles] = {};
};

// Build the fp16 image of W [N][Kin] (row-major f32) with rows taken in the order perm[] (size
// n_valid; nullptr = identity), padded to the N-tiles `tiles` and to Kp columns.
void build_layer(Layer &L, const float *W, const float *bias, const int *perm, int n_valid, int Kin, int Kp,
                 const std::vector<int> &tiles, bool pair = false) {
    L.K = Kp; L.n_valid = n_valid; L.n_tiles = (int)tiles.size();
    int col = 0; size_t off = 0;
    for (size_t j = 0; j < tiles.size(); ++j) {
        L.nt_size[j] = tiles[j]; L.nt_col[j] = col; L.nt_off[j] = (unsigned)off;
        col += tiles[j]; off += (size_t)tiles[j] * Kp * 2;
    }
    L.Nc = col;
    std::vector<__half> img(off / 2, __float2half(0.0f));
    std::vector<float> b((size_t)L.Nc, 0.0f);
    for (size_t j = 0; j < tiles.size(); ++j) {
        const int NT = tiles[j];
        uint8_t *base = reinterpret_cast<uint8_t *>(img.data()) + L.nt_off[j];
        for (int n = 0; n < NT; ++n) {
            const int p = L.nt_col[j] + n;
            if (p >= n_valid) continue;
            const int src = perm ? perm[p] : p;
            if (bias) b[(size_t)p] = bias[src];
            for (int k = 0; k < Kin; ++k)
                *reinterpret_cast<__half *>(base + tc::tiled_off(n, k, NT)) = __float2half_rn(W[(size_t)src * Kin + k]);
        }
    }
    L.img.ensure(img.size());
    FVAD_CUDA_CHECK(cudaMemcpy(L.img.p, img.data(), img.size() * sizeof(__half), cudaMemcpyHostToDevice));
    // CTA-pair image: same N-tile offsets, each tile split into two halves of NT/2 rows tiled with R = NT/2
    std::vector<__half> img2(off / 2, __float2half(0.0f));
    for (size_t j = 0; j < tiles.size(); ++j) {
        const int NT = tiles[j], NH = NT / 2;
        uint8_t *base = reinterpret_cast<uint8_t *>(img2.data()) + L.nt_off[j];
        for (int n = 0; n < NT; ++n) {
            const int p = L.nt_col[j] + n;
            if (p >= n_valid) continue;
            const int src = perm ? perm[p] : p;
            uint8_t *hb = base + (size_t)(n / NH) * NH * Kp * 2;
            for (int k = 0; k < Kin; ++k)
                *reinterpret_cast<__half *>(hb + tc::tiled_off(n % NH, k, NH)) = __float2half_rn(W[(size_t)src * Kin + k]);
        }
    }
    L.img2.ensure(img2.size());
    FVAD_CUDA_CHECK(cudaMemcpy(L.img2.p, img2.data(), img2.size() * sizeof(__half), cudaMemcpyHostToDevice));
    L.bias.ensure(b.size());
    FVAD_CUDA_CHECK(cudaMemcpy(L.bias.p, b.data(), b.size() * sizeof(float), cudaMemcpyHostToDevice));
    memset(&L.maps, 0, sizeof(L.maps));
    L.pair = pair;
    if (pair) tc::build_pair_maps(L.maps, L.img2.p, Kp, L.n_tiles, L.nt_size, L.nt_off);
}

class NnTc final : public NnBackend {
public:
    NnTc(const HostWeights &w, bool split3) {
        if (split3) throw Error(FVAD_ERR_INVALID_ARGUMENT, "FVAD_NN_TC_F16X3 is not built yet; use FVAD_NN_F32 for the 1e-4 gate");
        if (w.hidden != 400 || w.n_bins != 161 || w.fc2 != 600 || w.fc3 != 600)
            throw Error(FVAD_ERR_MODEL, "the tensor-core backend is specialised for the NSNet2-baseline shapes (161/400/600/600)");
        int dev = 0;
        FVAD_CUDA_CHECK(cudaGetDevice(&dev));
        FVAD_CUDA_CHECK(cudaDeviceGetAttribute(&n_sm_, cudaDevAttrMultiProcessorCount, dev));
        fc1_relu_ = w.fc1_relu;
        if (const char *e = getenv("FVAD_GRU_PAIR")) pair_ = atoi(e) != 0;
        if (const char *e = getenv("FVAD_NN_TIMING")) timing_ = atoi(e) != 0;
        if (const char *e = getenv("FVAD_GEMM_PAIR")) gemm_pair_ = atoi(e) != 0;
        if (const char *e = getenv("FVAD_GEMM_TS")) gemm_ts_ = atoi(e) != 0;
        // N-tiles: with the A tile in TMEM (K/2 columns) the two accumulators share what is left of the 512 columns
        const bool ts = gemm_ts_;
        using T = std::vector<int>;
        const int H = 400;
        // gate-row permutation: p = 48c + 16g + u  <-  g*400 + 16c + u
        std::vector<int> perm(3 * H);
        for (int c = 0; c < 25; ++c)
            for (int g = 0; g < 3; ++g)
                for (int u = 0; u < 16; ++u) perm[(size_t)(48 * c + 16 * g + u)] = g * H + 16 * c + u;
        // which layers run on CTA pairs (FVAD_GEMM_PAIR_MASK, A/B and bisecting): 1 fc1, 2 W1/W2, 4 fc2, 8 fc3, 16 fc4
        int pmask = gemm_pair_ ? 30 : 0;  // (fc1: K = 176 makes it accumulator-drain bound either way; one CTA is as fast)
        if (const char *e = getenv("FVAD_GEMM_PAIR_MASK")) pmask = atoi(e);
        bool gp = (pmask & 1) != 0;
        // CTA pairs (nn_tc_gemm2.cu): every layer's A tile in TMEM (K/2 columns); N-tiles of 96 leave room for three
        // accumulators at K = 400 (the layers whose accumulator drain runs as fast as their MMAs); the short tile comes
        // FIRST so that the M-pair's last N-tile is a long one and the next tile's A slabs restage under it
        build_layer(fc1_, w.fc1_w.data(), w.fc1_b.data(), nullptr, H, 161, 176, T{208, 192}, gp);
        for (int l = 0; l < 2; ++l) {
            // xp bias: Wb + Rb for z and r; Wb alone for n (Rb_n is applied inside r*(.), nn_tc_gru.cu)
            // The z and r rows of W, R and their biases are stored HALVED: the gates are sigma(x) = 0.5 + 0.5 tanh(x/2) on the
            // hardware tanh, and with x/2 coming out of the MMAs directly the epilogue saves a multiply per gate.  Exact: a
            // power of two commutes with every rounding on the way (fp16 weights, fp32 accumulation, fp16 projections), short
            // of fp16 subnormals.
            std::vector<float> b(3 * H), Wh(w.gru_W[l]);
            for (int i = 0; i < 3 * H; ++i) b[(size_t)i] = w.gru_Wb[l][(size_t)i] + (i < 2 * H ? w.gru_Rb[l][(size_t)i] : 0.0f);
            for (int i = 0; i < 2 * H; ++i) {
                b[(size_t)i] *= 0.5f;
                for (int k = 0; k < H; ++k) Wh[(size_t)i * H + k] *= 0.5f;
            }
            auto Rh = [&](int src, int k) { return (src < 2 * H ? 0.5f : 1.0f) * w.gru_R[l][(size_t)src * H + k]; };
            gp = (pmask & 2) != 0;
            build_layer(gW_[l], Wh.data(), b.data(), perm.data(), 3 * H, H, H,
                        gp ? T{48, 144, 144, 144, 144, 144, 144, 144, 144}
                           : ts ? T{144, 144, 144, 144, 144, 144, 144, 144, 48} : T{240, 240, 240, 240, 240}, gp);
            // R: 25 chunk images of 48 rows = one "layer" with 25 N-tiles of 48
            Layer tmp;
            std::vector<__half> img((size_t)3 * H * H);
            for (int p = 0; p < 3 * H; ++p) {
                const int c = p / 48, n = p % 48, src = perm[(size_t)p];
                uint8_t *base = reinterpret_cast<uint8_t *>(img.data()) + (size_t)c * 48 * H * 2;
                for (int k = 0; k < H; ++k)
                    *reinterpret_cast<__half *>(base + tc::tiled_off(n, k, 48)) = __float2half_rn(Rh(src, k));
            }
            gR_[l].ensure(img.size());
            FVAD_CUDA_CHECK(cudaMemcpy(gR_[l].p, img.data(), img.size() * sizeof(__half), cudaMemcpyHostToDevice));
            // CTA-pair image (nn_tc_gru2.cu): chunk c = [rows 0..23 | rows 24..47], each half tiled with R = 24
            std::vector<__half> img2((size_t)3 * H * H);
            for (int p = 0; p < 3 * H; ++p) {
                const int c = p / 48, n = p % 48, half = n / 24, src = perm[(size_t)p];
                uint8_t *base = reinterpret_cast<uint8_t *>(img2.data()) + ((size_t)c * 48 + (size_t)half * 24) * H * 2;
                for (int k = 0; k < H; ++k)
                    *reinterpret_cast<__half *>(base + tc::tiled_off(n % 24, k, 24)) = __float2half_rn(Rh(src, k));
            }
            gR2_[l].ensure(img2.size());
            FVAD_CUDA_CHECK(cudaMemcpy(gR2_[l].p, img2.data(), img2.size() * sizeof(__half), cudaMemcpyHostToDevice));
            tc::build_gru_pair_map(gRmap_[l], gR2_[l].p);
            gRbn_[l].ensure((size_t)H);
            FVAD_CUDA_CHECK(cudaMemcpy(gRbn_[l].p, w.gru_Rb[l].data() + 2 * H, H * sizeof(float), cudaMemcpyHostToDevice));
        }
        // fc1 folded into the first GRU's input projection.  When the graph has no activation between them (fc1_relu == 0:
        // the NSNet2 baseline), W1 (fc1 x + b1) + b = (W1 fc1) x + (W1 b1 + b): one K = 161 contraction instead of 161 -> 400
        // followed by 400 -> 1200, the same function with 333 000 fewer multiply-adds per frame-step (12 % of the network) and
        // one fp16 rounding of the activations less.  The product is formed in f64 and rounded to fp16 once.
        // FVAD_NN_FOLD_FC1=0 keeps the two layers apart (A/B; what the f32 backend and the oracle always do).
        fold_fc1_ = !w.fc1_relu;
        if (const char *e = getenv("FVAD_NN_FOLD_FC1")) fold_fc1_ = fold_fc1_ && atoi(e) != 0;
        if (fold_fc1_) {
            const int NB = 161;
            std::vector<float> Wf((size_t)3 * H * NB), bf((size_t)3 * H);
            for (int o = 0; o < 3 * H; ++o) {
                const float *wr = w.gru_W[0].data() + (size_t)o * H;
                double acc_b = (double)w.gru_Wb[0][(size_t)o] + (o < 2 * H ? (double)w.gru_Rb[0][(size_t)o] : 0.0);
                for (int k = 0; k < H; ++k) acc_b += (double)wr[k] * (double)w.fc1_b[(size_t)k];
                const double gs = o < 2 * H ? 0.5 : 1.0;  // z and r rows halved, see above
                bf[(size_t)o] = (float)(gs * acc_b);
                for (int i = 0; i < NB; ++i) {
                    double acc = 0.0;
                    for (int k = 0; k < H; ++k) acc += (double)wr[k] * (double)w.fc1_w[(size_t)k * NB + i];
                    Wf[(size_t)o * NB + i] = (float)(gs * acc);
                }
            }
            gp = (pmask & 2) != 0;
            // K = 176: the A tile takes 88 TMEM columns, which leaves room for three accumulators of 128 (the layer is
            // accumulator-drain bound: 1 200 fp32 columns to read back per 176-deep contraction)
            build_layer(gWf_, Wf.data(), bf.data(), perm.data(), 3 * H, NB, 176,
                        gp ? T{48, 128, 128, 128, 128, 128, 128, 128, 128, 128} : T{144, 144, 144, 144, 144, 144, 144, 144, 48}, gp);
            gWf_.a_tmem = ts ? 1 : 0;
            // Images of the fused first layer (nn_tc_gru3.cu): per chunk c two halves of 24 rows tiled with R = 24.
            //   R: rows [n z r] of units 16c..16c+15 (z, r halved), K = 400;
            //   W: rows [z r n] of the folded projection (z, r halved), K = 176, bias as fp16 hi + lo in columns 161, 162.
            fuse_ = pair_;
            if (const char *e = getenv("FVAD_GRU_FUSE")) fuse_ = fuse_ && atoi(e) != 0;
            if (fuse_) {
                std::vector<__half> rimg((size_t)3 * H * H, __float2half(0.0f)), wimg((size_t)3 * H * 176, __float2half(0.0f));
                for (int c = 0; c < 25; ++c)
                    for (int n = 0; n < 48; ++n) {
                        const int u = n % 16, half = n / 24, nr = n % 24;
                        {   // R row n of the chunk: gate order n, z, r
                            const int gate = n / 16 == 0 ? 2 : n / 16 - 1;  // ONNX gate index: z 0, r 1, n 2
                            const int src = gate * H + 16 * c + u;
                            uint8_t *base = reinterpret_cast<uint8_t *>(rimg.data()) + ((size_t)c * 48 + (size_t)half * 24) * H * 2;
                            const float gs = gate < 2 ? 0.5f : 1.0f;
                            for (int k = 0; k < H; ++k)
                                *reinterpret_cast<__half *>(base + tc::tiled_off(nr, k, 24)) = __float2half_rn(gs * w.gru_R[0][(size_t)src * H + k]);
                        }
                        {   // W row n of the chunk: gate order z, r, n (Wf / bf carry the halving already)
                            const int src = (n / 16) * H + 16 * c + u;
                            uint8_t *base = reinterpret_cast<uint8_t *>(wimg.data()) + ((size_t)c * 48 + (size_t)half * 24) * 176 * 2;
                            for (int k = 0; k < NB; ++k)
                                *reinterpret_cast<__half *>(base + tc::tiled_off(nr, k, 24)) = __float2half_rn(Wf[(size_t)src * NB + k]);
                            const __half hi = __float2half_rn(bf[(size_t)src]);
                            *reinterpret_cast<__half *>(base + tc::tiled_off(nr, 161, 24)) = hi;
                            *reinterpret_cast<__half *>(base + tc::tiled_off(nr, 162, 24)) = __float2half_rn(bf[(size_t)src] - __half2float(hi));
                        }
                    }
                gR3_.ensure(rimg.size());
                gW3_.ensure(wimg.size());
                FVAD_CUDA_CHECK(cudaMemcpy(gR3_.p, rimg.data(), rimg.size() * sizeof(__half), cudaMemcpyHostToDevice));
                FVAD_CUDA_CHECK(cudaMemcpy(gW3_.p, wimg.data(), wimg.size() * sizeof(__half), cudaMemcpyHostToDevice));
                tc::build_gru3_maps(gR3map_, gW3map_, gR3_.p, gW3_.p);
            }
        }
        gp = (pmask & 4) != 0;
        build_layer(fc2_, w.fc2_w.data(), w.fc2_b.data(), nullptr, 600, H, H, gp ? T{112, 112, 128, 128, 128} : T{208, 208, 192}, gp);
        gp = (pmask & 8) != 0;
        build_layer(fc3_, w.fc3_w.data(), w.fc3_b.data(), nullptr, 600, 600, 608, gp ? T{32, 96, 96, 96, 96, 96, 96} : T{208, 208, 192}, gp);
        gp = (pmask & 16) != 0;
        build_layer(fc4_, w.fc4_w.data(), w.fc4_b.data(), nullptr, 161, 600, 608, gp ? T{80, 96} : ts ? T{96, 80} : T{176}, gp);
        // A in TMEM caps the N-tile at (512 - K/2)/2 columns: 144 at K = 400, 96 at K = 608.  Measured per 20 stream-hours:
        // fc1 2.12 -> 1.75 ms, W1/W2 8.6 -> 7.5, fc4 3.43 -> 3.2; fc2 unchanged (4.0) and fc3 (K = 608, N-tile 96) 7.3 -> 8.8,
        // so those two keep A in shared memory with N-tiles of 208
        fc1_.a_tmem = gW_[0].a_tmem = gW_[1].a_tmem = fc4_.a_tmem = ts ? 1 : 0;
    }
    const char *name() const override { return "tcgen05-f16"; }

    bool wants_feat16() const override { return true; }

    void forward(const NnSpan &sp, cudaStream_t s) override {
        const int total = sp.SC * sp.K;
        const int q_total = (total + tc::kTileM - 1) / tc::kTileM;
        const int QB = std::min(q_total, 2 * n_sm_);  // tiles per sub-batch
        if (!sp.feat16) throw Error(FVAD_ERR_INVALID_ARGUMENT, "the tensor-core backend reads the fp16 feature rows");
        if (const char *e = getenv("FVAD_NN_GATHER")) gather_ = atoi(e) != 0;
        a0_.ensure((size_t)QB * kSteps * tc::tile_bytes(176) / 2);
        if (!fold_fc1_) x1_.ensure((size_t)QB * kSteps * tc::tile_bytes(400) / 2);
        xp_.ensure((size_t)QB * kSteps * tc::tile_bytes(1200) / 2);
        h1_.ensure((size_t)QB * kSteps * tc::tile_bytes(400) / 2);
        h2_.ensure((size_t)QB * kSteps * tc::tile_bytes(400) / 2);
        f2_.ensure((size_t)QB * kFrames * tc::tile_bytes(608) / 2);
        f3_.ensure((size_t)QB * kFrames * tc::tile_bytes(608) / 2);
        for (int q0 = 0; q0 < q_total; q0 += QB) {
            const int Q = std::min(QB, q_total - q0);
            const __half *a0 = a0_.p;
            tc::PackArgs pa{static_cast<const __half *>(sp.feat16), a0_.p, Q, q0 * tc::kTileM, total, sp.K};
            // the folded first projection on CTA pairs gathers its 54-row windows from the feature rows itself
            const bool fused = fold_fc1_ && fuse_;  // first GRU layer with its input projection fused in: no pack, no projection GEMM
            const bool gather = !fused && fold_fc1_ && gWf_.pair && gather_;
            mark(0, s);
            if (!gather && !fused) tc::launch_tc_pack(pa, s);
            mark(1, s);
            if (!fold_fc1_) gemm(fc1_, a0, x1_.p, Q, kSteps, kSteps, 0, kSteps, 0, fc1_relu_ ? 1 : 0, s);
            mark(2, s);
            if (fused) {
            } else if (gather) {
                GemmArgs g = base_args(gWf_, nullptr, xp_.p, Q, kSteps, kSteps, 0, kSteps, 0, 0);
                g.gather_rows = static_cast<const __half *>(sp.feat16); g.seq0 = q0 * tc::kTileM; g.n_seq = total; g.Kchunks = sp.K;
                maybe_trace(g, gWf_, s);
                launch(g, gWf_, s);
            } else if (fold_fc1_) gemm(gWf_, a0, xp_.p, Q, kSteps, kSteps, 0, kSteps, 0, 0, s);
            else gemm(gW_[0], x1_.p, xp_.p, Q, kSteps, kSteps, 0, kSteps, 0, 0, s);
            mark(3, s);
            if (fused) {
                tc::Gru3Args ga{static_cast<const __half *>(sp.feat16), gRbn_[0].p, h1_.p, Q, 0, q0 * tc::kTileM, total, sp.K};
                tc::launch_tc_gru3(ga, gR3map_, gW3map_, n_sm_, s);
            } else
            gru(0, h1_.p, Q, 0, s);
            mark(4, s);
            gemm(gW_[1], h1_.p, xp_.p, Q, kSteps, kSteps, 0, kSteps, 0, 0, s);
            mark(5, s);
            gru(1, h2_.p, Q, kCtx, s);
            mark(6, s);
            if (sp.after_rnn && q0 + QB >= q_total) FVAD_CUDA_CHECK(cudaEventRecord(sp.after_rnn, s));
            gemm(fc2_, h2_.p, f2_.p, Q, kFrames, kSteps, kCtx, kFrames, 0, 1, s);
            mark(7, s);
            gemm(fc3_, f2_.p, f3_.p, Q, kFrames, kFrames, 0, kFrames, 0, 1, s);
            mark(8, s);
            GemmArgs g = base_args(fc4_, f3_.p, nullptr, Q, kFrames, kFrames, 0, kFrames, 0, 2);
            g.out_mode = 1; g.gains = sp.gains; g.seq0 = q0 * tc::kTileM; g.n_seq = total; g.Kchunks = sp.K;
            launch(g, fc4_, s);
            mark(9, s);
            ++batches_;
        }
    }

    ~NnTc() override {
        if (timing_) {
            static const char *nm[9] = {"pack", "fc1", "W1", "gru1", "W2", "gru2", "fc2", "fc3", "fc4"};
            double tot = 0;
            for (double v : acc_ms_) tot += v;
            fprintf(stderr, "[fvad nn timing] total %.2f ms:", tot);
            for (int i = 0; i < 9; ++i) fprintf(stderr, " %s %.2f (%.0f%%)", nm[i], acc_ms_[i], 100.0 * acc_ms_[i] / (tot > 0 ? tot : 1));
            fprintf(stderr, "\n");
        }
        for (auto &e : tev_) cudaEventDestroy(e);
    }

    // per-layer CUDA-event time and launch counts since the last call (the caller has synchronised the stream)
    void take_layer_ms(double *ms9, unsigned *launches9) override {
        static const int per_batch[9] = {1, 1, 1, 1, 1, 1, 1, 1, 1};
        for (int i = 0; i < 9; ++i) { ms9[i] = 0; launches9[i] = 0; }
        for (size_t b = 0; b < batches_; ++b)
            for (int i = 0; i < 9; ++i) {
                float ms = 0;
                cudaEventElapsedTime(&ms, tev_[b * 10 + i], tev_[b * 10 + i + 1]);
                ms9[i] += ms; acc_ms_[i] += ms; launches9[i] += per_batch[i];
            }
        batches_ = 0;
    }

private:
    int n_sm_ = 148, fc1_relu_ = 0;
    Layer fc1_, gW_[2], gWf_, fc2_, fc3_, fc4_;
    bool fold_fc1_ = false;   // fc1 folded into W1 (gWf_): see the constructor
    bool fuse_ = false;       // ... and that projection fused into the first GRU kernel (nn_tc_gru3.cu; FVAD_GRU_FUSE=0: separate kernels)
    DevBuf<__half> gR3_, gW3_;
    CUtensorMap gR3map_, gW3map_;
    bool gather_ = false;     // FVAD_NN_GATHER=1: the folded projection's A stagers gather the 54-row windows from the feature rows
                              // themselves (no pack pass).  Parity-green, measured neutral per 20 stream-hours (pack 1.55 -> 0 ms, the
                              // projection 4.13 -> 5.67: thread-per-row 32-byte loads), so the pack kernel stays the default
    DevBuf<__half> gR_[2], gR2_[2];
    CUtensorMap gRmap_[2];  // TMA view of gR2_ (pair kernel)
    // per-layer CUDA-event timing (events only, no extra synchronisation); FVAD_NN_TIMING=1 prints the lifetime sums
    bool timing_ = false;
    std::vector<cudaEvent_t> tev_;   // 10 per sub-batch of the current push
    size_t batches_ = 0;
    double acc_ms_[9] = {};
    void mark(int i, cudaStream_t s) {
        const size_t k = batches_ * 10 + (size_t)i;
        while (tev_.size() <= k) { cudaEvent_t e; FVAD_CUDA_CHECK(cudaEventCreate(&e)); tev_.push_back(e); }
        cudaEventRecord(tev_[k], s);
    }
    DevBuf<unsigned long long> trace_, gtrace_;
    bool gtraced_ = false;
    bool traced_ = false;
    bool pair_ = true;        // GRU on CTA pairs (nn_tc_gru2.cu); FVAD_GRU_PAIR=0: the one-CTA kernel (A/B)
    bool gemm_ts_ = true;     // A tile in TMEM (FVAD_GEMM_TS=0: resident in shared memory, the first design)
    bool gemm_pair_ = true;   // CTA-pair GEMM (nn_tc_gemm2.cu) for W1/W2/fc2/fc3/fc4; FVAD_GEMM_PAIR=0: every layer on the one-CTA kernel (A/B)

    void gru(int l, __half *hout, int Q, int t_store_min, cudaStream_t s) {
        unsigned long long *tr = nullptr;
        const char *tp = getenv("FVAD_GRU_TRACE");
        if (tp && l == 0 && !traced_) { trace_.ensure(54 * 25 * 8); FVAD_CUDA_CHECK(cudaMemsetAsync(trace_.p, 0, 54 * 25 * 8 * 8, s)); tr = trace_.p; }
        if (pair_) {
            tc::GruArgs ga{xp_.p, gR2_[l].p, gRbn_[l].p, hout, Q, t_store_min, tr};
            if (const char *e = getenv("FVAD_GRU_DBG")) ga.dbg = atoi(e);
            tc::launch_tc_gru2(ga, gRmap_[l], n_sm_, s);
        }
        else tc::launch_tc_gru(tc::GruArgs{xp_.p, gR_[l].p, gRbn_[l].p, hout, Q, t_store_min, tr}, n_sm_, s);
        if (tr) {  // debug only: dump CTA 0's timeline of the first GRU launch
            traced_ = true;
            std::vector<unsigned long long> h(54 * 25 * 8);
            FVAD_CUDA_CHECK(cudaStreamSynchronize(s));
            FVAD_CUDA_CHECK(cudaMemcpy(h.data(), trace_.p, h.size() * 8, cudaMemcpyDeviceToHost));
            if (FILE *f = fopen(tp, "wb")) { fwrite(h.data(), 8, h.size(), f); fclose(f); }
        }
    }
    DevBuf<float> gRbn_[2];
    DevBuf<__half> a0_, x1_, xp_, h1_, h2_, f2_, f3_;

    static GemmArgs base_args(const Layer &L, const __half *A, __half *C, int Q, int T, int A_q, int A_t0, int C_q, int C_t0, int act) {
        GemmArgs g;
        memset(&g, 0, sizeof(g));
        g.A = A; g.B = L.img.p; g.B2 = L.img2.p; g.bias = L.bias.p; g.C = C;
        g.K = L.K; g.Nc = L.Nc; g.n_valid = L.n_valid; g.n_tiles = L.n_tiles;
        for (int j = 0; j < L.n_tiles; ++j) { g.nt_size[j] = L.nt_size[j]; g.nt_col[j] = L.nt_col[j]; g.nt_off[j] = L.nt_off[j]; }
        if (const char *e = getenv("FVAD_GEMM_DBG")) g.dbg = atoi(e);
        g.a_tmem = L.a_tmem;
        g.mt_count = Q * T; g.T = T; g.A_q = A_q; g.A_t0 = A_t0; g.C_q = C_q; g.C_t0 = C_t0; g.act = act;
        return g;
    }
    void launch(const GemmArgs &g, const Layer &L, cudaStream_t s) {
        if (L.pair) tc::launch_tc_gemm2(g, L.maps, n_sm_, s);
        else tc::launch_tc_gemm(g, n_sm_, s);
    }
    void gemm(const Layer &L, const __half *A, __half *C, int Q, int T, int A_q, int A_t0, int C_q, int C_t0, int act, cudaStream_t s) {
        GemmArgs g = base_args(L, A, C, Q, T, A_q, A_t0, C_q, C_t0, act);
        maybe_trace(g, L, s);
        launch(g, L, s);
    }
    // debug only (FVAD_GEMM_TRACE=<file> [FVAD_GEMM_TRACE_LAYER=fc1|W1|W2|fc2|fc3]): one extra launch of that layer with
    // CTA 0's timeline probe on, dumped to the file (tools/gemm_trace.py, tools/gemm_pair_trace.py)
    void maybe_trace(GemmArgs g, const Layer &L, cudaStream_t s) {
        const char *tp = getenv("FVAD_GEMM_TRACE");
        if (!tp || gtraced_) return;
        const char *want = getenv("FVAD_GEMM_TRACE_LAYER");
        const Layer *sel = fold_fc1_ ? &gWf_ : &gW_[0];
        if (want) sel = !strcmp(want, "fc1") ? &fc1_ : !strcmp(want, "W2") ? &gW_[1] : !strcmp(want, "fc2") ? &fc2_ : !strcmp(want, "fc3") ? &fc3_ : sel;
        if (sel != &L) return;
        gtraced_ = true;
        const size_t ne = 5 * 4 * 8 * 32;
        gtrace_.ensure(ne);
        FVAD_CUDA_CHECK(cudaMemsetAsync(gtrace_.p, 0, ne * 8, s));
        g.trace = gtrace_.p;
        launch(g, L, s);
        std::vector<unsigned long long> h(ne);
        FVAD_CUDA_CHECK(cudaStreamSynchronize(s));
        FVAD_CUDA_CHECK(cudaMemcpy(h.data(), gtrace_.p, ne * 8, cudaMemcpyDeviceToHost));
        if (FILE *f = fopen(tp, "wb")) { fwrite(h.data(), 8, ne, f); fclose(f); }
    }
};

}  // namespace

std::unique_ptr<NnBackend> make_nn_tc(const HostWeights &w, bool split3) { return std::make_unique<NnTc>(w, split3); }

}  // namespace fvad
