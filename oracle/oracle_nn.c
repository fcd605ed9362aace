/*
 * oracle_nn.c -- CPU ORACLE (test infrastructure only; see fvad_oracle.h).
 *
 * Restates what `onnx_instance.run()` (src/NSNet2.zig:220) computes for the NSNet2 baseline
 * graph: FC 161->400, GRU 400 x2, FC 400->600 ReLU, FC 600->600 ReLU, FC 600->161 sigmoid
 * (SURVEY App. C), on a [steps,161] input with ZERO initial hidden state on every call.
 * GRU follows the ONNX operator definition with linear_before_reset = 1 and gate order z,r,h:
 *   z = s(Wz x + Rz h + Wbz + Rbz);  r = s(Wr x + Rr h + Wbr + Rbr)
 *   n = tanh(Wh x + Wbh + r * (Rh h + Rbh));  h' = (1-z)*n + z*h
 * Third-party arithmetic being restated: onnxruntime 1.15.0 CPU EP (prebuilt, absent from the
 * snapshot) -- PARITY UNPINNED; cross-checked against torch.nn.GRU in tests/test_oracle_cpu.py.
 */
#include "fvad_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

struct orc_nn {
    int nb, H, F2, F3, fc1_relu;
    /* weights stored transposed [in][out] so the inner loops run over contiguous outputs */
    float *fc1_wt, *fc1_b;
    float *g_Wt[2], *g_Rt[2], *g_Wb[2], *g_Rb[2];
    float *fc2_wt, *fc2_b, *fc3_wt, *fc3_b, *fc4_wt, *fc4_b;
    /* scratch, sized on first use */
    int cap_steps;
    float *a0, *a1, *gx, *gh;
};

static float *transpose_copy(const float *w, int out, int in) {
    float *t = (float *)malloc(sizeof(float) * (size_t)out * (size_t)in);
    for (int o = 0; o < out; ++o)
        for (int i = 0; i < in; ++i) t[(size_t)i * out + o] = w[(size_t)o * in + i];
    return t;
}
static float *dup(const float *v, int n) {
    float *d = (float *)malloc(sizeof(float) * (size_t)n);
    memcpy(d, v, sizeof(float) * (size_t)n);
    return d;
}

orc_nn *orc_nn_new(const fvad_nsnet2_weights *w) {
    orc_nn *p = (orc_nn *)calloc(1, sizeof(*p));
    p->nb = w->n_bins; p->H = w->hidden; p->F2 = w->fc2; p->F3 = w->fc3; p->fc1_relu = w->fc1_relu;
    const int H = p->H;
    p->fc1_wt = transpose_copy(w->fc1_w, H, p->nb);
    p->fc1_b = dup(w->fc1_b, H);
    const float *W[2] = {w->gru1_W, w->gru2_W}, *R[2] = {w->gru1_R, w->gru2_R};
    const float *Wb[2] = {w->gru1_Wb, w->gru2_Wb}, *Rb[2] = {w->gru1_Rb, w->gru2_Rb};
    for (int l = 0; l < 2; ++l) {
        p->g_Wt[l] = transpose_copy(W[l], 3 * H, H);
        p->g_Rt[l] = transpose_copy(R[l], 3 * H, H);
        p->g_Wb[l] = dup(Wb[l], 3 * H);
        p->g_Rb[l] = dup(Rb[l], 3 * H);
    }
    p->fc2_wt = transpose_copy(w->fc2_w, p->F2, H);   p->fc2_b = dup(w->fc2_b, p->F2);
    p->fc3_wt = transpose_copy(w->fc3_w, p->F3, p->F2); p->fc3_b = dup(w->fc3_b, p->F3);
    p->fc4_wt = transpose_copy(w->fc4_w, p->nb, p->F3); p->fc4_b = dup(w->fc4_b, p->nb);
    return p;
}

void orc_nn_free(orc_nn *p) {
    if (!p) return;
    free(p->fc1_wt); free(p->fc1_b);
    for (int l = 0; l < 2; ++l) { free(p->g_Wt[l]); free(p->g_Rt[l]); free(p->g_Wb[l]); free(p->g_Rb[l]); }
    free(p->fc2_wt); free(p->fc2_b); free(p->fc3_wt); free(p->fc3_b); free(p->fc4_wt); free(p->fc4_b);
    free(p->a0); free(p->a1); free(p->gx); free(p->gh);
    free(p);
}

/* C[M][N] = A[M][K] * Wt[K][N] + b ; row-blocked so each weight row is reused across 6 rows */
static void affine(const float *A, const float *Wt, const float *b, float *C, int M, int K, int N) {
    enum { RB = 6 };
    for (int m0 = 0; m0 < M; m0 += RB) {
        const int mb = (M - m0 < RB) ? (M - m0) : RB;
        for (int r = 0; r < mb; ++r) memcpy(C + (size_t)(m0 + r) * N, b, sizeof(float) * (size_t)N);
        for (int k = 0; k < K; ++k) {
            const float *wr = Wt + (size_t)k * N;
            for (int r = 0; r < mb; ++r) {
                const float a = A[(size_t)(m0 + r) * K + k];
                float *c = C + (size_t)(m0 + r) * N;
                for (int n = 0; n < N; ++n) c[n] += a * wr[n];
            }
        }
    }
}

static inline float sigm(float x) { return 1.0f / (1.0f + expf(-x)); }

/* one GRU layer over `steps` rows; in/out [steps][H]; zero initial state */
static void gru_layer(orc_nn *p, int l, const float *in, float *out, int steps) {
    const int H = p->H, G = 3 * H;
    affine(in, p->g_Wt[l], p->g_Wb[l], p->gx, steps, H, G); /* W x + Wb for every step */
    float *h = (float *)calloc((size_t)H, sizeof(float));
    for (int t = 0; t < steps; ++t) {
        affine(h, p->g_Rt[l], p->g_Rb[l], p->gh, 1, H, G); /* R h + Rb */
        const float *gx = p->gx + (size_t)t * G, *gh = p->gh;
        float *o = out + (size_t)t * H;
        for (int j = 0; j < H; ++j) {
            const float z = sigm(gx[j] + gh[j]);
            const float r = sigm(gx[H + j] + gh[H + j]);
            const float n = tanhf(gx[2 * H + j] + r * gh[2 * H + j]);
            o[j] = (1.0f - z) * n + z * h[j];
        }
        memcpy(h, o, sizeof(float) * (size_t)H);
    }
    free(h);
}

void orc_nn_forward(orc_nn *p, const float *features, float *gains, int steps) {
    const int H = p->H;
    if (steps > p->cap_steps) {
        free(p->a0); free(p->a1); free(p->gx); free(p->gh);
        const size_t wide = (size_t)(p->F2 > p->F3 ? p->F2 : p->F3);
        const size_t w = wide > (size_t)H ? wide : (size_t)H;
        p->a0 = (float *)malloc(sizeof(float) * (size_t)steps * w);
        p->a1 = (float *)malloc(sizeof(float) * (size_t)steps * w);
        p->gx = (float *)malloc(sizeof(float) * (size_t)steps * 3 * (size_t)H);
        p->gh = (float *)malloc(sizeof(float) * 3 * (size_t)H);
        p->cap_steps = steps;
    }
    affine(features, p->fc1_wt, p->fc1_b, p->a0, steps, p->nb, H);
    if (p->fc1_relu)
        for (int i = 0; i < steps * H; ++i) p->a0[i] = p->a0[i] > 0 ? p->a0[i] : 0;
    gru_layer(p, 0, p->a0, p->a1, steps);
    gru_layer(p, 1, p->a1, p->a0, steps);
    affine(p->a0, p->fc2_wt, p->fc2_b, p->a1, steps, H, p->F2);
    for (int i = 0; i < steps * p->F2; ++i) p->a1[i] = p->a1[i] > 0 ? p->a1[i] : 0;
    affine(p->a1, p->fc3_wt, p->fc3_b, p->a0, steps, p->F2, p->F3);
    for (int i = 0; i < steps * p->F3; ++i) p->a0[i] = p->a0[i] > 0 ? p->a0[i] : 0;
    affine(p->a0, p->fc4_wt, p->fc4_b, gains, steps, p->F3, p->nb);
    for (int i = 0; i < steps * p->nb; ++i) gains[i] = sigm(gains[i]);
}

/* ---- double-precision evaluation of the same graph (tolerance studies) ---- */

static void affine64(const double *A, const float *W /*[N][K]*/, const float *b, double *C, int M, int K, int N) {
    for (int m = 0; m < M; ++m)
        for (int n = 0; n < N; ++n) {
            double acc = b[n];
            const float *w = W + (size_t)n * K;
            const double *a = A + (size_t)m * K;
            for (int k = 0; k < K; ++k) acc += a[k] * (double)w[k];
            C[(size_t)m * N + n] = acc;
        }
}
static double sigm64(double x) { return 1.0 / (1.0 + exp(-x)); }

static void gru64(const float *W, const float *R, const float *Wb, const float *Rb, const double *in, double *out, int steps, int H) {
    const int G = 3 * H;
    double *gx = (double *)malloc(sizeof(double) * (size_t)steps * G);
    double *gh = (double *)malloc(sizeof(double) * (size_t)G);
    double *h = (double *)calloc((size_t)H, sizeof(double));
    affine64(in, W, Wb, gx, steps, H, G);
    for (int t = 0; t < steps; ++t) {
        affine64(h, R, Rb, gh, 1, H, G);
        const double *x = gx + (size_t)t * G;
        for (int j = 0; j < H; ++j) {
            const double z = sigm64(x[j] + gh[j]);
            const double r = sigm64(x[H + j] + gh[H + j]);
            const double n = tanh(x[2 * H + j] + r * gh[2 * H + j]);
            out[(size_t)t * H + j] = (1.0 - z) * n + z * h[j];
        }
        memcpy(h, out + (size_t)t * H, sizeof(double) * (size_t)H);
    }
    free(gx); free(gh); free(h);
}

void orc_nn_forward_f64(const fvad_nsnet2_weights *w, const float *features, double *gains, int steps) {
    const int H = w->hidden, nb = w->n_bins;
    const size_t wide = 1024;
    double *f = (double *)malloc(sizeof(double) * (size_t)steps * wide);
    double *a = (double *)malloc(sizeof(double) * (size_t)steps * wide);
    double *b = (double *)malloc(sizeof(double) * (size_t)steps * wide);
    for (int i = 0; i < steps * nb; ++i) f[i] = features[i];
    affine64(f, w->fc1_w, w->fc1_b, a, steps, nb, H);
    if (w->fc1_relu) for (int i = 0; i < steps * H; ++i) a[i] = a[i] > 0 ? a[i] : 0;
    gru64(w->gru1_W, w->gru1_R, w->gru1_Wb, w->gru1_Rb, a, b, steps, H);
    gru64(w->gru2_W, w->gru2_R, w->gru2_Wb, w->gru2_Rb, b, a, steps, H);
    affine64(a, w->fc2_w, w->fc2_b, b, steps, H, w->fc2);
    for (int i = 0; i < steps * w->fc2; ++i) b[i] = b[i] > 0 ? b[i] : 0;
    affine64(b, w->fc3_w, w->fc3_b, a, steps, w->fc2, w->fc3);
    for (int i = 0; i < steps * w->fc3; ++i) a[i] = a[i] > 0 ? a[i] : 0;
    affine64(a, w->fc4_w, w->fc4_b, gains, steps, w->fc3, nb);
    for (int i = 0; i < steps * nb; ++i) gains[i] = sigm64(gains[i]);
    free(f); free(a); free(b);
}
