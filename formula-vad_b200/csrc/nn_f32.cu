// nn_f32.cu -- NSNet2 in plain fp32 on the CUDA cores.  This is the arithmetic of the reference
// (onnxruntime CPU EP computes in f32) and the parity anchor for the tensor-core backends; it is
// a correct-first implementation (64x64 register-tiled GEMMs), not the fast path.
#include <algorithm>

#include "nn.cuh"

namespace fvad {

namespace {

constexpr int TM = 64, TN = 64, TK = 16;

struct RowMap {  // how a GEMM row index maps to memory
    int mode;    // 0 dense: base + r*ld ; 1 feature window (read) ; 2 gains scatter (write)
    int Bq, q0, K;  // sequences in this sub-batch, first global sequence, chunks per stream-channel
    int t0;         // time offset added to the row's step
};

__device__ __forceinline__ size_t map_row(const RowMap &m, int r, int ld) {
    if (m.mode == 0) return (size_t)r * ld;
    const int t = r / m.Bq + m.t0, q = r % m.Bq;
    const int g = m.q0 + q, sc = g / m.K, j = g % m.K;
    if (m.mode == 1) return ((size_t)sc * (kCtx + kFrames * m.K) + (size_t)kFrames * j + t) * ld;
    return ((size_t)sc * (kFrames * m.K) + (size_t)kFrames * j + t) * ld;  // mode 2, t already excludes ctx
}

// C = act(A * W^T + b).  A rows via amap, C rows via cmap.  W [N][K] row-major.  act: 0 none, 1 relu, 2 sigmoid
__global__ void __launch_bounds__(256) gemm_f32_kernel(const float *__restrict__ A, int lda, RowMap amap,
                                                       const float *__restrict__ W, const float *__restrict__ bias,
                                                       float *__restrict__ C, int ldc, RowMap cmap, int M, int N, int K, int act) {
    __shared__ float As[TK][TM + 1];
    __shared__ float Ws[TK][TN + 1];
    const int tid = threadIdx.x, tx = tid % 16, ty = tid / 16;
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < K; k0 += TK) {
        for (int i = tid; i < TM * TK; i += 256) {
            const int r = i / TK, k = i % TK;
            const int gr = m0 + r, gk = k0 + k;
            As[k][r] = (gr < M && gk < K) ? A[map_row(amap, gr, lda) + gk] : 0.0f;
        }
        for (int i = tid; i < TN * TK; i += 256) {
            const int r = i / TK, k = i % TK;
            const int gn = n0 + r, gk = k0 + k;
            Ws[k][r] = (gn < N && gk < K) ? W[(size_t)gn * K + gk] : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < TK; ++k) {
            float a[4], w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[k][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) w[j] = Ws[k][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gr = m0 + ty * 4 + i;
        if (gr >= M) continue;
        float *crow = C + map_row(cmap, gr, ldc);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn >= N) continue;
            float v = acc[i][j] + bias[gn];
            if (act == 1) v = fmaxf(v, 0.0f);
            else if (act == 2) v = 1.0f / (1.0f + expf(-v));
            crow[gn] = v;
        }
    }
}

// One GRU time step for Bq sequences (ONNX GRU, linear_before_reset=1, gate order z,r,h):
//   gh = h_prev * R^T + Rb ; z = s(xz+ghz) ; r = s(xr+ghr) ; n = tanh(xn + r*ghn) ; h = (1-z)*n + z*h_prev
// xp = W x + Wb is precomputed for all steps.  Tile: 64 sequences x 32 hidden units (96 gate rows of R).
constexpr int GU = 32;
__global__ void __launch_bounds__(256) gru_step_f32_kernel(const float *__restrict__ xp, const float *__restrict__ hprev,
                                                           const float *__restrict__ R, const float *__restrict__ Rb,
                                                           float *__restrict__ hout, int Bq, int H, int first) {
    __shared__ float Hs[TK][TM + 1];
    __shared__ float Rs[3][TK][GU + 1];
    const int tid = threadIdx.x, tx = tid % 16, ty = tid / 16;
    const int m0 = blockIdx.y * TM, u0 = blockIdx.x * GU;
    float acc[3][4][2] = {};
    if (!first) {
        for (int k0 = 0; k0 < H; k0 += TK) {
            for (int i = tid; i < TM * TK; i += 256) {
                const int r = i / TK, k = i % TK;
                Hs[k][r] = (m0 + r < Bq) ? hprev[(size_t)(m0 + r) * H + k0 + k] : 0.0f;
            }
            for (int i = tid; i < 3 * GU * TK; i += 256) {
                const int g = i / (GU * TK), rem = i % (GU * TK), u = rem / TK, k = rem % TK;
                Rs[g][k][u] = (u0 + u < H) ? R[((size_t)g * H + u0 + u) * H + k0 + k] : 0.0f;
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < TK; ++k) {
                float a[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) a[i] = Hs[k][ty * 4 + i];
#pragma unroll
                for (int g = 0; g < 3; ++g)
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const float w = Rs[g][k][tx * 2 + j];
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[g][i][j] = fmaf(a[i], w, acc[g][i][j]);
                    }
            }
            __syncthreads();
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int q = m0 + ty * 4 + i;
        if (q >= Bq) continue;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int u = u0 + tx * 2 + j;
            if (u >= H) continue;
            const float *x = xp + (size_t)q * 3 * H;
            const float ghz = acc[0][i][j] + Rb[u], ghr = acc[1][i][j] + Rb[H + u], ghn = acc[2][i][j] + Rb[2 * H + u];
            const float z = 1.0f / (1.0f + expf(-(x[u] + ghz)));
            const float r = 1.0f / (1.0f + expf(-(x[H + u] + ghr)));
            const float n = tanhf(x[2 * H + u] + r * ghn);
            const float hp = first ? 0.0f : hprev[(size_t)q * H + u];
            hout[(size_t)q * H + u] = (1.0f - z) * n + z * hp;
        }
    }
}

class NnF32 final : public NnBackend {
public:
    explicit NnF32(const HostWeights &w) : H(w.hidden), B(w.n_bins), F2(w.fc2), F3(w.fc3), fc1_relu(w.fc1_relu) {
        up(fc1_w, w.fc1_w); up(fc1_b, w.fc1_b);
        for (int l = 0; l < 2; ++l) { up(gW[l], w.gru_W[l]); up(gR[l], w.gru_R[l]); up(gWb[l], w.gru_Wb[l]); up(gRb[l], w.gru_Rb[l]); }
        up(fc2_w, w.fc2_w); up(fc2_b, w.fc2_b); up(fc3_w, w.fc3_w); up(fc3_b, w.fc3_b); up(fc4_w, w.fc4_w); up(fc4_b, w.fc4_b);
    }
    const char *name() const override { return "f32-simt"; }

    void forward(const NnSpan &sp, cudaStream_t s) override {
        const int total = sp.SC * sp.K;
        const int BQ = std::min(total, 4096);
        x1.ensure((size_t)kSteps * BQ * H);
        xp.ensure((size_t)kSteps * BQ * 3 * H);
        g1.ensure((size_t)kSteps * BQ * H);
        f2.ensure((size_t)kFrames * BQ * std::max(F2, F3));
        f3.ensure((size_t)kFrames * BQ * std::max(F2, F3));
        for (int q0 = 0; q0 < total; q0 += BQ) {
            const int Bq = std::min(BQ, total - q0);
            const RowMap dense{0, Bq, q0, sp.K, 0};
            const RowMap win{1, Bq, q0, sp.K, 0};
            const RowMap scat{2, Bq, q0, sp.K, 0};
            const int M = kSteps * Bq, M50 = kFrames * Bq;
            gemm(sp.feat, B, win, fc1_w.p, fc1_b.p, x1.p, H, dense, M, H, B, fc1_relu ? 1 : 0, s);
            gru(0, x1.p, g1.p, Bq, s);
            gru(1, g1.p, x1.p, Bq, s);  // layer-2 output reuses x1
            const float *g2 = x1.p + (size_t)kCtx * Bq * H;  // rows of steps 4..53
            gemm(g2, H, dense, fc2_w.p, fc2_b.p, f2.p, F2, dense, M50, F2, H, 1, s);
            gemm(f2.p, F2, dense, fc3_w.p, fc3_b.p, f3.p, F3, dense, M50, F3, F2, 1, s);
            gemm(f3.p, F3, dense, fc4_w.p, fc4_b.p, sp.gains, B, scat, M50, B, F3, 2, s);
        }
        if (sp.after_rnn) FVAD_CUDA_CHECK(cudaEventRecord(sp.after_rnn, s));
    }

private:
    int H, B, F2, F3, fc1_relu;
    DevBuf<float> fc1_w, fc1_b, gW[2], gR[2], gWb[2], gRb[2], fc2_w, fc2_b, fc3_w, fc3_b, fc4_w, fc4_b;
    DevBuf<float> x1, xp, g1, f2, f3;

    static void up(DevBuf<float> &d, const std::vector<float> &h) {
        d.ensure(h.size());
        FVAD_CUDA_CHECK(cudaMemcpy(d.p, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    static void gemm(const float *A, int lda, RowMap am, const float *W, const float *b, float *C, int ldc, RowMap cm,
                     int M, int N, int K, int act, cudaStream_t s) {
        dim3 grid((unsigned)((N + TN - 1) / TN), (unsigned)((M + TM - 1) / TM));
        gemm_f32_kernel<<<grid, 256, 0, s>>>(A, lda, am, W, b, C, ldc, cm, M, N, K, act);
        FVAD_LAUNCHED();
    }
    void gru(int l, const float *in, float *out, int Bq, cudaStream_t s) {
        const RowMap dense{0, Bq, 0, 1, 0};
        gemm(in, H, dense, gW[l].p, gWb[l].p, xp.p, 3 * H, dense, kSteps * Bq, 3 * H, H, 0, s);
        dim3 grid((unsigned)((H + GU - 1) / GU), (unsigned)((Bq + TM - 1) / TM));
        for (int t = 0; t < kSteps; ++t) {
            const float *hp = t ? out + (size_t)(t - 1) * Bq * H : nullptr;
            gru_step_f32_kernel<<<grid, 256, 0, s>>>(xp.p + (size_t)t * Bq * 3 * H, hp, gR[l].p, gRb[l].p,
                                                     out + (size_t)t * Bq * H, Bq, H, t == 0);
            FVAD_LAUNCHED();
        }
        FVAD_CUDA_CHECK(cudaGetLastError());
    }
};

}  // namespace

std::unique_ptr<NnBackend> make_nn_f32(const HostWeights &w) { return std::make_unique<NnF32>(w); }

}  // namespace fvad
