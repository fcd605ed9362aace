/*
 * oracle_vad.c -- CPU ORACLE (test infrastructure only; see fvad_oracle.h).
 * Restates src/structures/RollingAverage.zig, src/AudioPipeline/VADMetadata.zig and
 * src/AudioPipeline/VADMachine.zig line by line (f64 rolling means, f32 elsewhere).
 */
#include "fvad_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ RollingAverage.zig */

struct orc_rolling {
    double *data;
    size_t len, write_idx, written;
    int has_last;
    double last_avg;
};

/* RollingAverage.zig:45-56: array-order sum of data[i] * (1/written), recomputed in full */
static double rolling_avg(orc_rolling *p) {
    double acc = 0.0;
    const double scalar = 1.0 / (double)p->written;
    for (size_t i = 0; i < p->written; ++i) acc += p->data[i] * scalar;
    p->last_avg = acc;
    p->has_last = 1;
    return acc;
}

/* RollingAverage.zig:11-28 */
orc_rolling *orc_rolling_new(size_t count, int has_init, double init) {
    orc_rolling *p = (orc_rolling *)calloc(1, sizeof(*p));
    p->data = (double *)calloc(count ? count : 1, sizeof(double));
    p->len = count;
    if (has_init) {
        for (size_t i = 0; i < count; ++i) p->data[i] = init;
        p->written = count;
        rolling_avg(p);
    }
    return p;
}
void orc_rolling_free(orc_rolling *p) {
    if (!p) return;
    free(p->data);
    free(p);
}
/* RollingAverage.zig:34-43 */
double orc_rolling_push(orc_rolling *p, float v) {
    p->data[p->write_idx] = v;
    p->write_idx = (p->write_idx + 1) % p->len;
    if (p->written < p->len) p->written += 1;
    return rolling_avg(p);
}
int orc_rolling_last(const orc_rolling *p, double *out) {
    if (p->has_last && out) *out = p->last_avg;
    return p->has_last;
}

/* ------------------------------------------------------------------ VADMetadata.zig */

/* VADMetadata.zig:29-60 */
void orc_meta_push(orc_meta *m, const orc_meta_result *v, float weight) {
    if (v->has_ratio) {
        if (!m->has_ratio) {
            m->has_ratio = 1;
            m->ratio_sum = 0.0f;
            m->ratio_weight = 0.0f;
        }
        m->ratio_sum += v->ratio * weight;
        m->ratio_weight += weight;
    }
    if (v->has_min && (!m->has_min || v->vmin < m->vmin)) { m->has_min = 1; m->vmin = v->vmin; }
    if (v->has_max && (!m->has_max || v->vmax > m->vmax)) { m->has_max = 1; m->vmax = v->vmax; }
}

/* VADMetadata.zig:16-27 */
orc_meta_result orc_meta_to_result(const orc_meta *m) {
    orc_meta_result r;
    memset(&r, 0, sizeof(r));
    r.has_min = m->has_min; r.vmin = m->vmin;
    r.has_max = m->has_max; r.vmax = m->vmax;
    if (m->has_ratio) { r.has_ratio = 1; r.ratio = m->ratio_sum / m->ratio_weight; }
    return r;
}

/* ------------------------------------------------------------------ VADMachine.zig */

enum { ST_CLOSED = 0, ST_OPENING = 1, ST_OPEN = 2, ST_CLOSING = 3 };

struct orc_vadm {
    fvad_vad_machine_config cfg;
    size_t sample_rate, n_channels, fft_size;
    int state;
    orc_rolling *long_term, *short_term, *ratio;
    int has_start, has_end;
    uint64_t start_index, end_index;
    float ratio_sum; size_t ratio_count; float met_sec;
    fvad_segment *segs; size_t n_segs, cap_segs;
    int last_met;
};

/* VADMachine.zig:75-136 */
orc_vadm *orc_vadm_new(const fvad_vad_machine_config *cfg, size_t sample_rate, size_t n_channels, size_t fft_size) {
    orc_vadm *p = (orc_vadm *)calloc(1, sizeof(*p));
    p->cfg = *cfg;
    p->sample_rate = sample_rate; p->n_channels = n_channels; p->fft_size = fft_size;
    const float evals = (float)sample_rate / (float)fft_size;
    size_t n_long = (size_t)(evals * cfg->long_term_speech_avg_sec);
    size_t n_short = (size_t)(evals * cfg->short_term_speech_avg_sec);
    size_t n_ratio = (size_t)(evals * cfg->channel_vol_ratio_avg_sec);
    p->long_term = orc_rolling_new(n_long > 1 ? n_long : 1, cfg->has_initial_long_term_avg, cfg->initial_long_term_avg);
    p->short_term = orc_rolling_new(n_short > 1 ? n_short : 1, 0, 0);
    p->ratio = orc_rolling_new(n_ratio, 0, 0); /* no @max(1, ..) in the reference (:102-106) */
    p->state = ST_CLOSED;
    return p;
}
void orc_vadm_free(orc_vadm *p) {
    if (!p) return;
    orc_rolling_free(p->long_term); orc_rolling_free(p->short_term); orc_rolling_free(p->ratio);
    free(p->segs);
    free(p);
}
size_t orc_vadm_segments(const orc_vadm *p, const fvad_segment **segs) { if (segs) *segs = p->segs; return p->n_segs; }
int orc_vadm_last_met(const orc_vadm *p) { return p->last_met; }
int orc_vadm_state(const orc_vadm *p) { return p->state; }

/* VADMachine.zig:312-325 */
static uint64_t rec_start(const orc_vadm *p, uint64_t from) {
    const uint64_t margin = (uint64_t)((float)p->sample_rate * 2);
    return from - (margin < from ? margin : from);
}
static uint64_t rec_end(const orc_vadm *p, uint64_t to) {
    return to + (uint64_t)((float)p->sample_rate * 2);
}

/* VADMachine.zig:265-309 */
static int on_speech_end(orc_vadm *p, uint64_t *sample_number) {
    const float sr = (float)p->sample_rate;
    const uint64_t from = p->start_index, to = p->end_index;
    const float len_sec = (float)(to - from) / sr;
    const int long_enough = len_sec >= p->cfg.min_vad_duration_sec;
    const float avg_ratio = p->ratio_sum / (float)p->ratio_count;
    if (long_enough) {
        if (p->n_segs == p->cap_segs) {
            p->cap_segs = p->cap_segs ? p->cap_segs * 2 : 100;
            p->segs = (fvad_segment *)realloc(p->segs, sizeof(fvad_segment) * p->cap_segs);
        }
        fvad_segment s;
        s.sample_from = rec_start(p, from);
        s.sample_to = rec_end(p, to);
        s.avg_channel_vol_ratio = avg_ratio;
        s.vad_met_sec = p->met_sec;
        p->segs[p->n_segs++] = s;
        *sample_number = rec_end(p, to);
        return 2; /* completed */
    }
    *sample_number = 0;
    return 3; /* aborted */
}

/* VADMachine.zig:138-263 (run + trackSpeechStats).  channel_band = averageVolumeInBand output. */
int orc_vadm_run(orc_vadm *p, const float *channel_band, uint64_t index, int has_ratio, float ratio_in, uint64_t *sample_number) {
    const float sr = (float)p->sample_rate;
    const fvad_vad_machine_config *c = &p->cfg;
    float min_volume = 999, max_volume = 0;
    for (size_t ch = 0; ch < p->n_channels; ++ch) {
        if (channel_band[ch] < min_volume) min_volume = channel_band[ch];
        if (channel_band[ch] > max_volume) max_volume = channel_band[ch];
    }
    (void)max_volume;
    const uint64_t min_to_open = (uint64_t)(sr * c->min_consecutive_sec_to_open);
    const uint64_t max_gap = (uint64_t)(sr * c->max_speech_gap_sec);
    const float meta_ratio = has_ratio ? ratio_in : 0.0f;

    const double short_term = orc_rolling_push(p->short_term, min_volume);
    const double vol_ratio = orc_rolling_push(p->ratio, meta_ratio);

    double base;
    if (!orc_rolling_last(p->long_term, &base)) base = c->has_initial_long_term_avg ? c->initial_long_term_avg : short_term;
    const double threshold = base * (double)c->speech_threshold_factor;
    const int met = (short_term > threshold) && (vol_ratio > (double)c->channel_vol_ratio_threshold);
    p->last_met = met;
    if (!met) orc_rolling_push(p->long_term, min_volume);

    int rec = 0;
    uint64_t num = 0;
    const int from_state = p->state;
    switch (p->state) {
    case ST_CLOSED:
        if (met) { p->state = ST_OPENING; p->start_index = index; p->has_start = 1; }
        break;
    case ST_OPENING: {
        const int dur_ok = (index - p->start_index) >= min_to_open;
        if (met && dur_ok) { p->state = ST_OPEN; rec = 1; num = rec_start(p, p->start_index); }
        else if (!met) p->state = ST_CLOSED;
        break;
    }
    case ST_OPEN:
        if (!met) { p->state = ST_CLOSING; p->end_index = index; p->has_end = 1; }
        break;
    case ST_CLOSING: {
        const int dur_ok = (index - p->end_index) >= max_gap;
        if (met) p->state = ST_OPEN;
        else if (dur_ok) { p->state = ST_CLOSED; rec = on_speech_end(p, &num); }
        break;
    }
    }
    /* trackSpeechStats :241-263 */
    const float frame_sec = (float)p->fft_size / sr;
    if (from_state == ST_CLOSED && p->state == ST_OPENING) {
        p->ratio_sum = meta_ratio;
        p->ratio_count = 1;
        p->met_sec = frame_sec;
    } else if (from_state == ST_OPEN) {
        p->ratio_sum += meta_ratio;
        p->ratio_count += 1;
        if (met) p->met_sec += frame_sec;
    }
    if (sample_number) *sample_number = num;
    return rec;
}
