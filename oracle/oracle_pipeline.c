/*
 * oracle_pipeline.c -- CPU ORACLE (test infrastructure only; see fvad_oracle.h).
 * Restates the step order of src/AudioPipeline.zig:118-143 (pushSamples) ->
 * src/AudioPipeline/VADPipeline.zig:140-237 (collectInputStep / denoiserStep / fftStep /
 * stateMachineStep) with BufferedVolumeAnalyzer.zig:29-69, BufferedDenoiser.zig:75-120,
 * BufferedFFT.zig:74-202 and the re-blocking semantics of SegmentWriter.zig:46-114.
 * The ring buffers / recorders are host plumbing and are not restated: chunks are cut at
 * absolute offsets k*chunk exactly as VADPipeline.zig:150-153 does.
 */
#include "fvad_oracle.h"

#include <alloca.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct grow_f32 { float *p; size_t n, cap; } grow_f32;
static void gf_push(grow_f32 *g, const float *v, size_t n) {
    if (g->n + n > g->cap) {
        size_t c = g->cap ? g->cap : 4096;
        while (c < g->n + n) c *= 2;
        g->p = (float *)realloc(g->p, c * sizeof(float));
        g->cap = c;
    }
    memcpy(g->p + g->n, v, n * sizeof(float));
    g->n += n;
}

struct orc_pipeline {
    fvad_config cfg;
    size_t sr, nch, chunk, fft_size;
    uint64_t total_written, read_count;
    /* unread input per channel (stands in for the original_audio_buffer ring) */
    float **pend; size_t n_pend, cap_pend;
    orc_nsnet2 **den;        /* one per channel, BufferedDenoiser.zig:38-41 */
    float **den_out;         /* [nch][chunk] temp_result_segment */
    /* BufferedFFT */
    orc_fftop *fft; float *window; float norm; fvad_complex *cbuf;
    float **fbuf; size_t f_fill; uint64_t f_index; orc_meta f_meta;
    float **mag;             /* [nch][bins] */
    /* machines: [0] main, [1..] alt */
    size_t n_mach; orc_vadm **mach; int *bin_lo, *bin_hi;
    float *band_tmp;
    /* taps */
    grow_f32 t_band, t_ratio; uint8_t *t_met; size_t n_frames, cap_frames;
    int keep_den; grow_f32 *t_den;
    int keep_nn; grow_f32 t_feat, t_gain, t_spec;
    const float *forced; size_t forced_frames; size_t chunk_no;
};

orc_pipeline *orc_pipeline_new(const fvad_config *cfg, const fvad_nsnet2_weights *w, int allow_any_rate, int *err) {
    int e = 0;
    if (cfg->sample_rate != 48000 && !(allow_any_rate && cfg->sample_rate == 16000)) e = -1; /* VADPipeline.zig:55-58 */
    if (!e && (cfg->vad_config.fft_size == 0 || (cfg->vad_config.fft_size & 1))) e = -2;
    if (e) { if (err) *err = e; return NULL; }
    orc_pipeline *p = (orc_pipeline *)calloc(1, sizeof(*p));
    p->cfg = *cfg;
    p->sr = cfg->sample_rate; p->nch = cfg->n_channels; p->fft_size = cfg->vad_config.fft_size;
    p->chunk = orc_nsnet2_chunk_size(p->sr);
    const size_t C = p->nch, N = p->fft_size, bins = N / 2 + 1;
    p->pend = (float **)calloc(C, sizeof(float *));
    p->den = (orc_nsnet2 **)calloc(C, sizeof(*p->den));
    p->den_out = (float **)calloc(C, sizeof(float *));
    p->fbuf = (float **)calloc(C, sizeof(float *));
    p->mag = (float **)calloc(C, sizeof(float *));
    p->t_den = (grow_f32 *)calloc(C, sizeof(grow_f32));
    for (size_t c = 0; c < C; ++c) {
        p->den[c] = orc_nsnet2_new(p->sr, w);
        p->den_out[c] = (float *)calloc(p->chunk, sizeof(float));
        p->fbuf[c] = (float *)calloc(N, sizeof(float));
        p->mag[c] = (float *)calloc(bins, sizeof(float));
    }
    /* BufferedFFT.init :74-119 */
    p->fft = orc_fftop_new((int)N, 0);
    p->window = (float *)malloc(sizeof(float) * N);
    orc_hann_periodic(p->window, (int)N);
    p->norm = orc_window_norm_factor(p->window, (int)N) / (float)N;
    p->cbuf = (fvad_complex *)malloc(sizeof(fvad_complex) * bins);
    /* machines, VADPipeline.zig:109-123 */
    p->n_mach = 1 + (cfg->vad_config.alt_vad_machine_configs ? cfg->vad_config.n_alt_vad_machine_configs : 0);
    p->mach = (orc_vadm **)calloc(p->n_mach, sizeof(*p->mach));
    p->bin_lo = (int *)calloc(p->n_mach, sizeof(int));
    p->bin_hi = (int *)calloc(p->n_mach, sizeof(int));
    for (size_t m = 0; m < p->n_mach; ++m) {
        const fvad_vad_machine_config *mc = m ? &cfg->vad_config.alt_vad_machine_configs[m - 1] : &cfg->vad_config.vad_machine_config;
        p->mach[m] = orc_vadm_new(mc, p->sr, C, N);
        p->bin_lo[m] = orc_freq_to_bin(mc->speech_min_freq, p->sr, N);
        p->bin_hi[m] = orc_freq_to_bin(mc->speech_max_freq, p->sr, N);
    }
    p->band_tmp = (float *)calloc(C, sizeof(float));
    if (err) *err = 0;
    return p;
}

void orc_pipeline_free(orc_pipeline *p) {
    if (!p) return;
    for (size_t c = 0; c < p->nch; ++c) {
        free(p->pend[c]); orc_nsnet2_free(p->den[c]); free(p->den_out[c]); free(p->fbuf[c]); free(p->mag[c]); free(p->t_den[c].p);
    }
    free(p->pend); free(p->den); free(p->den_out); free(p->fbuf); free(p->mag); free(p->t_den);
    orc_fftop_free(p->fft); free(p->window); free(p->cbuf);
    for (size_t m = 0; m < p->n_mach; ++m) orc_vadm_free(p->mach[m]);
    free(p->mach); free(p->bin_lo); free(p->bin_hi); free(p->band_tmp);
    free(p->t_band.p); free(p->t_ratio.p); free(p->t_met);
    free(p->t_feat.p); free(p->t_gain.p); free(p->t_spec.p);
    free(p);
}

uint64_t orc_pipeline_total_write_count(const orc_pipeline *p) { return p->total_written; }
size_t orc_pipeline_segments(const orc_pipeline *p, size_t m, const fvad_segment **segs) {
    if (m >= p->n_mach) { if (segs) *segs = NULL; return 0; }
    return orc_vadm_segments(p->mach[m], segs);
}
size_t orc_pipeline_frames(const orc_pipeline *p, const float **band, const float **ratio, const uint8_t **met) {
    if (band) *band = p->t_band.p;
    if (ratio) *ratio = p->t_ratio.p;
    if (met) *met = p->t_met;
    return p->n_frames;
}
void orc_pipeline_keep_denoised(orc_pipeline *p, int on) { p->keep_den = on; }
size_t orc_pipeline_denoised(const orc_pipeline *p, size_t ch, const float **pcm) {
    if (pcm) *pcm = p->t_den[ch].p;
    return p->t_den[ch].n;
}
void orc_pipeline_keep_nn_taps(orc_pipeline *p, int on) { p->keep_nn = on; }
size_t orc_pipeline_nn_taps(const orc_pipeline *p, const float **f, const float **g, const fvad_complex **s) {
    if (f) *f = p->t_feat.p;
    if (g) *g = p->t_gain.p;
    if (s) *s = (const fvad_complex *)p->t_spec.p;
    return p->t_feat.n / FVAD_NS_BINS;
}
void orc_pipeline_force_gains(orc_pipeline *p, const float *gains, size_t n_frames) { p->forced = gains; p->forced_frames = n_frames; }

/* stateMachineStep, VADPipeline.zig:209-237 (recording start/stop side effects are out of scope) */
static void state_machine_step(orc_pipeline *p, uint64_t index, const orc_meta_result *meta) {
    for (size_t m = 0; m < p->n_mach; ++m) {
        /* averageVolumeInBand, BufferedFFT.zig:183-202: a sequential f32 SUM over the band */
        for (size_t c = 0; c < p->nch; ++c) {
            float s = 0.0f;
            for (int b = p->bin_lo[m]; b < p->bin_hi[m] + 1; ++b) s += p->mag[c][b];
            p->band_tmp[c] = s;
        }
        uint64_t num;
        orc_vadm_run(p->mach[m], p->band_tmp, index, meta->has_ratio, meta->ratio, &num);
        if (m == 0) {
            float mn = 999;
            for (size_t c = 0; c < p->nch; ++c) if (p->band_tmp[c] < mn) mn = p->band_tmp[c];
            float r = meta->has_ratio ? meta->ratio : 0.0f;
            gf_push(&p->t_band, &mn, 1);
            gf_push(&p->t_ratio, &r, 1);
            if (p->n_frames == p->cap_frames) {
                p->cap_frames = p->cap_frames ? p->cap_frames * 2 : 4096;
                p->t_met = (uint8_t *)realloc(p->t_met, p->cap_frames);
            }
            p->t_met[p->n_frames++] = (uint8_t)orc_vadm_last_met(p->mach[0]);
        }
    }
}

/* fftStep + BufferedFFT.write/fft, VADPipeline.zig:191-207, BufferedFFT.zig:129-181 */
static void fft_step(orc_pipeline *p, uint64_t seg_index, const orc_meta_result *in_meta) {
    const size_t N = p->fft_size, bins = N / 2 + 1;
    size_t off = 0;
    while (off < p->chunk) {
        size_t room = N - p->f_fill, left = p->chunk - off;
        size_t n = room < left ? room : left;
        for (size_t c = 0; c < p->nch; ++c) memcpy(p->fbuf[c] + p->f_fill, p->den_out[c] + off, n * sizeof(float));
        p->f_fill += n;
        orc_meta_push(&p->f_meta, in_meta, (float)n);
        off += n;
        if (p->f_fill < N) return;
        for (size_t c = 0; c < p->nch; ++c) {
            orc_fftop_forward(p->fft, p->fbuf[c], p->window, p->cbuf);
            for (size_t b = 0; b < bins; ++b)
                p->mag[c][b] = sqrtf(p->cbuf[b].r * p->cbuf[b].r + p->cbuf[b].i * p->cbuf[b].i) * p->norm;
        }
        orc_meta_result fm = orc_meta_to_result(&p->f_meta);
        memset(&p->f_meta, 0, sizeof(p->f_meta));
        state_machine_step(p, p->f_index, &fm);
        p->f_fill = 0;
        p->f_index = seg_index + off; /* buffer.reset(input.segment.index + input_offset + n_written) */
    }
}

/* one chunk: collectInputStep body + denoiserStep */
static int process_chunk(orc_pipeline *p, uint64_t from) {
    const size_t C = p->nch, L = p->chunk;
    /* BufferedVolumeAnalyzer.analyseVolume :48-69 */
    float vmin = 1, vmax = 0;
    for (size_t c = 0; c < C; ++c) {
        float v = orc_rms_volume(p->pend[c], L);
        if (v < vmin) vmin = v;
        if (v > vmax) vmax = v;
    }
    orc_meta_result va; memset(&va, 0, sizeof(va));
    va.has_ratio = va.has_min = va.has_max = 1;
    va.ratio = (vmax == 0) ? 0.0f : vmin / vmax; va.vmin = vmin; va.vmax = vmax;
    orc_meta m1; memset(&m1, 0, sizeof(m1));
    orc_meta_push(&m1, &va, (float)L);                 /* BufferedVolumeAnalyzer.write :35-38 */
    orc_meta_result r1 = orc_meta_to_result(&m1);
    orc_meta m2; memset(&m2, 0, sizeof(m2));
    orc_meta_push(&m2, &r1, (float)L);                 /* BufferedDenoiser.write :83-86 */
    orc_meta_result r2 = orc_meta_to_result(&m2);
    for (size_t c = 0; c < C; ++c) {                   /* BufferedDenoiser.write :105-110 */
        if (p->forced) orc_nsnet2_force_gains(p->den[c], p->forced + ((size_t)c * p->forced_frames + p->chunk_no * FVAD_NS_FRAMES) * FVAD_NS_BINS);
        int rc = orc_nsnet2_denoise(p->den[c], p->pend[c], L, p->den_out[c]);
        if (rc) return rc;
        if (p->keep_den) gf_push(&p->t_den[c], p->den_out[c], L);
    }
    if (p->keep_nn) {
        const size_t cur = (size_t)FVAD_NS_CTX * FVAD_NS_BINS, nn = (size_t)FVAD_NS_FRAMES * FVAD_NS_BINS;
        gf_push(&p->t_feat, orc_nsnet2_last_features(p->den[0]) + cur, nn);
        gf_push(&p->t_gain, orc_nsnet2_last_gains(p->den[0]) + cur, nn);
        gf_push(&p->t_spec, (const float *)orc_nsnet2_last_spec(p->den[0]), nn * 2);
    }
    p->chunk_no++;
    fft_step(p, from, &r2);
    return 0;
}

/* AudioPipeline.pushSamples :118-143 + VADPipeline.collectInputStep :144-166 */
int orc_pipeline_push(orc_pipeline *p, const float *const *pcm, size_t n, uint64_t *first_index) {
    if (first_index) *first_index = p->total_written;
    if (p->n_pend + n > p->cap_pend) {
        size_t c = p->cap_pend ? p->cap_pend : p->chunk * 2;
        while (c < p->n_pend + n) c *= 2;
        for (size_t ch = 0; ch < p->nch; ++ch) p->pend[ch] = (float *)realloc(p->pend[ch], c * sizeof(float));
        p->cap_pend = c;
    }
    for (size_t ch = 0; ch < p->nch; ++ch) memcpy(p->pend[ch] + p->n_pend, pcm[ch], n * sizeof(float));
    p->n_pend += n;
    p->total_written += n;
    if (p->cfg.skip_processing) { p->n_pend = 0; return 0; }
    size_t used = 0;
    while (p->n_pend - used >= p->chunk) {
        /* shift the view instead of copying: process in place */
        float **save = (float **)alloca(sizeof(float *) * p->nch);
        for (size_t ch = 0; ch < p->nch; ++ch) { save[ch] = p->pend[ch]; p->pend[ch] += used; }
        int rc = process_chunk(p, p->read_count);
        for (size_t ch = 0; ch < p->nch; ++ch) p->pend[ch] = save[ch];
        if (rc) return rc;
        used += p->chunk;
        p->read_count += p->chunk;
    }
    if (used) {
        for (size_t ch = 0; ch < p->nch; ++ch) memmove(p->pend[ch], p->pend[ch] + used, (p->n_pend - used) * sizeof(float));
        p->n_pend -= used;
    }
    return 0;
}
