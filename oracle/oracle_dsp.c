/*
 * oracle_dsp.c -- CPU ORACLE (test infrastructure only; see fvad_oracle.h).
 * Restates: src/audio_utils/window_fn.zig, src/audio_utils/resample.zig, src/audio_utils.zig
 * (rmsVolume), src/FFT.zig (fft/invFft/freqToBin) and src/NSNet2.zig (the denoiser object).
 * f32 is used exactly where the reference uses f32.
 */
#include "fvad_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ windows */

/* window_fn.zig:51-68 with K=1, alphas {0.5, 0.5} (hannWindowPeriodic :22-28) */
void orc_hann_periodic(float *w, int n) {
    const float N = (float)n;
    const float two_pi = (float)(2 * 3.14159265358979323846);
    const float alpha[2] = {0.5f, 1.0f - 0.5f};
    for (int i = 0; i < n; ++i) {
        const float fi = (float)i;
        float acc = 0;
        for (int k = 0; k < 2; ++k) {
            const float fk = (float)k;
            const float sign = (k & 1) ? -1.0f : 1.0f; /* pow(-1, k) */
            acc += sign * alpha[k] * cosf((two_pi * fk * fi) / N);
        }
        w[i] = acc;
    }
}

/* window_fn.zig:30-41 */
void orc_hann_symmetric(float *w, int n) {
    const float N = (float)n;
    const float step = (float)(2 * 3.14159265358979323846) / (N - 1);
    for (int i = 0; i < n; ++i) w[i] = 0.5f - 0.5f * cosf((float)i * step);
}

/* window_fn.zig:8-16 */
float orc_window_norm_factor(const float *w, int n) {
    float sum = 0;
    for (int i = 0; i < n; ++i) sum += w[i];
    return (float)n / sum;
}

/* NSNet2.zig:384-396 */
void orc_nsnet2_window(float *w) {
    orc_hann_symmetric(w, FVAD_NS_NFFT);
    for (int i = 0; i < FVAD_NS_NFFT; ++i) w[i] = sqrtf(w[i]);
}

/* ------------------------------------------------------------------ small helpers */

/* audio_utils.zig:14-24: sequential f32 accumulation */
float orc_rms_volume(const float *x, size_t n) {
    float sum = 0.0f;
    for (size_t i = 0; i < n; ++i) sum += x[i] * x[i];
    return sqrtf(sum / (float)n);
}

/* resample.zig:9-29: bare decimation */
void orc_downsample(const float *in, size_t n_in, float *out, int rate) {
    const size_t steps = n_in / (size_t)rate;
    for (size_t i = 0; i < steps; ++i) out[i] = in[i * (size_t)rate];
}

/* resample.zig:67-79 (std.math.lerp is a fused multiply-add: a + (b-a)*t) */
static void interp(float a, float b, float *dst, int n) {
    for (int i = 0; i < n; ++i) {
        const float frac = (float)(i + 1) / (float)(n + 1);
        dst[i] = fmaf(b - a, frac, a);
    }
}

/* resample.zig:32-65 */
float orc_upsample(const float *in, size_t n_in, float *out, float prev_last, int rate) {
    const int gap = rate - 1;
    interp(prev_last, in[0], out, gap);
    out[gap] = in[0];
    float last = in[0];
    for (size_t i = 1; i < n_in; ++i) {
        float *dst = out + i * (size_t)rate;
        interp(in[i - 1], in[i], dst, gap);
        dst[gap] = in[i];
        last = in[i];
    }
    return last;
}

/* FFT.zig:140-167 */
int orc_freq_to_bin(float freq, size_t sample_rate, size_t n_fft) {
    const float nyq = (float)sample_rate / 2;
    if (freq > nyq) return -1;
    if (freq < 0) return -2;
    const float width = (float)sample_rate / (float)n_fft;
    return (int)roundf(freq / width);
}

/* ------------------------------------------------------------------ FFT operator */

static orc_kiss_alloc_fn g_kiss_alloc;
static orc_kiss_fftr_fn g_kiss_fwd;
static orc_kiss_fftri_fn g_kiss_inv;

void orc_use_kissfft(orc_kiss_alloc_fn alloc, orc_kiss_fftr_fn fwd, orc_kiss_fftri_fn inv) {
    g_kiss_alloc = alloc;
    g_kiss_fwd = fwd;
    g_kiss_inv = inv;
}

struct orc_fftop {
    int n;
    orc_rfft *own;  /* restated transform */
    void *kiss;     /* reference kissfft cfg when routed through oracle/_ref */
    float *buf;
};

orc_fftop *orc_fftop_new(int n_fft, int inverse) {
    if (n_fft <= 0 || (n_fft & 1)) return NULL; /* FFT.zig:41-43 InvalidFFTSize */
    orc_fftop *p = (orc_fftop *)calloc(1, sizeof(*p));
    p->n = n_fft;
    p->buf = (float *)malloc(sizeof(float) * (size_t)n_fft);
    if (g_kiss_alloc) p->kiss = g_kiss_alloc(n_fft, inverse, NULL, NULL);
    else p->own = orc_rfft_new(n_fft, inverse);
    return p;
}

void orc_fftop_free(orc_fftop *p) {
    if (!p) return;
    orc_rfft_free(p->own);
    free(p->kiss); /* kiss_fftr_free is plain free (kiss_fft.h) */
    free(p->buf);
    free(p);
}

/* FFT.zig:85-113 + loadSamplesFwd :183-199 */
void orc_fftop_forward(orc_fftop *p, const float *samples, const float *window, fvad_complex *bins) {
    for (int i = 0; i < p->n; ++i) p->buf[i] = samples[i] * window[i];
    if (p->kiss) g_kiss_fwd(p->kiss, p->buf, bins);
    else orc_rfft_forward(p->own, p->buf, bins);
}

/* FFT.zig:115-134 */
void orc_fftop_inverse(orc_fftop *p, const fvad_complex *bins, float *result) {
    if (p->kiss) g_kiss_inv(p->kiss, bins, result);
    else orc_rfft_inverse(p->own, bins, result);
}

/* ------------------------------------------------------------------ NSNet2 denoiser */

#define NFFT FVAD_NS_NFFT
#define HOP FVAD_NS_HOP
#define BINS FVAD_NS_BINS
#define FRAMES FVAD_NS_FRAMES
#define CTX FVAD_NS_CTX
#define STEPS FVAD_NS_STEPS
#define CHUNK FVAD_NS_CHUNK16

struct orc_nsnet2 {
    size_t in_rate;
    int down;
    orc_fftop *fwd, *inv;
    orc_nn *nn;
    float window[NFFT];
    float audio_in[CHUNK + HOP];
    float audio_out[CHUNK + HOP];
    fvad_complex spec[FRAMES * BINS];
    fvad_complex spec_pre[FRAMES * BINS]; /* tap: before gain */
    float ifft_buf[NFFT];
    float features[STEPS * BINS];
    float gains[STEPS * BINS];
    float last_sample;
    const float *forced; /* optional [50][161] */
};

size_t orc_nsnet2_chunk_size(size_t sample_rate) { return (size_t)CHUNK * (sample_rate / 16000); }

/* NSNet2.zig:35-141 */
orc_nsnet2 *orc_nsnet2_new(size_t sample_rate, const fvad_nsnet2_weights *w) {
    if (sample_rate % 16000 != 0) return NULL; /* resample.zig:4-7 panics */
    orc_nsnet2 *p = (orc_nsnet2 *)calloc(1, sizeof(*p));
    p->in_rate = sample_rate;
    p->down = (int)(sample_rate / 16000);
    p->fwd = orc_fftop_new(NFFT, 0);
    p->inv = orc_fftop_new(NFFT, 1);
    p->nn = w ? orc_nn_new(w) : NULL;
    orc_nsnet2_window(p->window);
    return p;
}

void orc_nsnet2_free(orc_nsnet2 *p) {
    if (!p) return;
    orc_fftop_free(p->fwd);
    orc_fftop_free(p->inv);
    orc_nn_free(p->nn);
    free(p);
}

void orc_nsnet2_force_gains(orc_nsnet2 *p, const float *g) { p->forced = g; }
const fvad_complex *orc_nsnet2_last_spec(const orc_nsnet2 *p) { return p->spec_pre; }
const float *orc_nsnet2_last_features(const orc_nsnet2 *p) { return p->features; }
const float *orc_nsnet2_last_gains(const orc_nsnet2 *p) { return p->gains; }

/* NSNet2.zig:161-237 */
int orc_nsnet2_denoise(orc_nsnet2 *p, const float *samples, size_t n, float *out) {
    if (n != (size_t)CHUNK * (size_t)p->down) return -1; /* InvalidInputLength :167-169 */
    const size_t cur = (size_t)CTX * BINS; /* features.len - n_frames*n_bins */

    /* :196-203 carries */
    memcpy(p->audio_in, p->audio_in + CHUNK, sizeof(float) * HOP);
    memcpy(p->audio_out, p->audio_out + CHUNK, sizeof(float) * HOP);
    memset(p->audio_out + HOP, 0, sizeof(float) * CHUNK);
    memmove(p->features, p->features + (size_t)FRAMES * BINS, sizeof(float) * cur);

    orc_downsample(samples, n, p->audio_in + HOP, p->down); /* :205-209 */

    /* calcSpectrogram :239-264 */
    for (int f = 0; f < FRAMES; ++f)
        orc_fftop_forward(p->fwd, p->audio_in + f * HOP, p->window, p->spec + f * BINS);
    memcpy(p->spec_pre, p->spec, sizeof(p->spec));

    /* calcFeatures :266-287 */
    const float p_min = powf(10.0f, -12.0f);
    for (int i = 0; i < FRAMES * BINS; ++i) {
        const fvad_complex b = p->spec[i];
        const float pw = b.r * b.r + b.i * b.i;
        p->features[cur + i] = log10f(pw > p_min ? pw : p_min);
    }

    /* onnx_instance.run() :220 */
    if (p->forced) {
        memset(p->gains, 0, sizeof(float) * cur);
        memcpy(p->gains + cur, p->forced, sizeof(float) * FRAMES * BINS);
    } else if (p->nn) {
        orc_nn_forward(p->nn, p->features, p->gains, STEPS);
    } else {
        return -2;
    }

    /* applySpecgramGain :289-310 */
    for (int i = 0; i < FRAMES * BINS; ++i) {
        float g = p->gains[cur + i];
        if (g < -80.0f) g = -80.0f;
        else if (g > 1.0f) g = 1.0f;
        p->spec[i].r *= g;
        p->spec[i].i *= g;
    }

    /* reconstructAudio :312-339 */
    const float norm = 1.0f / (float)NFFT;
    for (int f = 0; f < FRAMES; ++f) {
        orc_fftop_inverse(p->inv, p->spec + f * BINS, p->ifft_buf);
        for (int i = 0; i < NFFT; ++i) {
            p->ifft_buf[i] *= p->window[i] * norm;
            p->audio_out[f * HOP + i] += p->ifft_buf[i];
        }
    }

    /* :231-236 */
    p->last_sample = orc_upsample(p->audio_out, CHUNK, out, p->last_sample, p->down);
    return 0;
}
