/*
 * oracle_eval.c -- CPU ORACLE (test infrastructure only; see fvad_oracle.h).
 * Restates src/Evaluator.zig:90-156 (sort by start, O(n*m) overlap matching),
 * src/Evaluator/SpeechSegment.zig:18-56 and src/Evaluator/statistics.zig:85-114,170-284
 * (extrusion + gap filling, TP/FP/FN seconds, rates, F-score, Fowlkes-Mallows).  All f32.
 * Pinned by the reference's own KATs statistics.zig:286-360 (tests/test_oracle_kats.py).
 */
#include "fvad_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

static float seg_overlap(orc_eval_seg a, orc_eval_seg b) { /* SpeechSegment.zig:22-27 */
    const float f = a.from_sec > b.from_sec ? a.from_sec : b.from_sec;
    const float t = a.to_sec < b.to_sec ? a.to_sec : b.to_sec;
    return t - f;
}

static int by_start(const void *a, const void *b) {
    const float x = ((const orc_eval_seg *)a)->from_sec, y = ((const orc_eval_seg *)b)->from_sec;
    return (x > y) - (x < y);
}

/* statistics.zig:191-203 with extrudeSegments :229-256 and calcOverlapMany :279-283.
 * `refs` are the reference segments matched to this VAD segment, in start order. */
float orc_eval_false_positive_sec(const orc_eval_seg *vad, const orc_eval_seg *refs, size_t n, const orc_eval_cfg *cfg) {
    float overlap = 0.0f;
    if (n) {
        orc_eval_seg *c = (orc_eval_seg *)malloc(sizeof(*c) * n);
        memcpy(c, refs, sizeof(*c) * n);
        c[0].from_sec -= cfg->extrude_start;
        c[n - 1].to_sec += cfg->extrude_end;
        for (size_t i = 0; i + 1 < n; ++i)
            if (c[i + 1].from_sec - c[i].to_sec <= cfg->fill_gaps) c[i].to_sec = c[i + 1].from_sec;
        for (size_t i = 0; i < n; ++i) {
            const float o = seg_overlap(*vad, c[i]);
            overlap += o > 0.0f ? o : 0.0f;
        }
        free(c);
    }
    const float fp = (vad->to_sec - vad->from_sec) - overlap;
    return fp > 0.0f ? fp : 0.0f;
}

void orc_evaluate(const orc_eval_seg *vad_in, size_t n_vad, const orc_eval_seg *ref_in, size_t n_ref,
                  const orc_eval_cfg *cfg, orc_eval_stats *st) {
    memset(st, 0, sizeof(*st));
    orc_eval_seg *vad = (orc_eval_seg *)malloc(sizeof(*vad) * (n_vad ? n_vad : 1));
    orc_eval_seg *ref = (orc_eval_seg *)malloc(sizeof(*ref) * (n_ref ? n_ref : 1));
    orc_eval_seg *match = (orc_eval_seg *)malloc(sizeof(*match) * (n_ref ? n_ref : 1));
    memcpy(vad, vad_in, sizeof(*vad) * n_vad);
    memcpy(ref, ref_in, sizeof(*ref) * n_ref);
    /* Evaluator.zig:110-111 (std.mem.sort is stable; ties keep input order -> use a stable pass) */
    for (size_t i = 1; i < n_vad; ++i) { orc_eval_seg k = vad[i]; size_t j = i; while (j && by_start(&vad[j - 1], &k) > 0) { vad[j] = vad[j - 1]; --j; } vad[j] = k; }
    for (size_t i = 1; i < n_ref; ++i) { orc_eval_seg k = ref[i]; size_t j = i; while (j && by_start(&ref[j - 1], &k) > 0) { ref[j] = ref[j - 1]; --j; } ref[j] = k; }

    /* statistics.zig:88-94 */
    for (size_t i = 0; i < n_vad; ++i) {
        size_t nm = 0;
        for (size_t j = 0; j < n_ref; ++j)
            if (seg_overlap(vad[i], ref[j]) > 0.0f) match[nm++] = ref[j]; /* findOverlapping */
        const float fp = orc_eval_false_positive_sec(&vad[i], match, nm, cfg);
        float tp = (vad[i].to_sec - vad[i].from_sec) - fp; /* calcTruePositiveSec :205-214 */
        if (tp < 0.0f) tp = 0.0f;
        st->fp_sec += fp;
        st->tp_sec += tp;
        st->p_sec += tp;
    }
    /* statistics.zig:96-102 */
    for (size_t j = 0; j < n_ref; ++j) {
        const float dur = ref[j].to_sec - ref[j].from_sec;
        if (dur < cfg->ignore_shorter_than) continue;
        float ov = 0.0f;
        for (size_t i = 0; i < n_vad; ++i) {
            const float o = seg_overlap(ref[j], vad[i]);
            if (o > 0.0f) ov += o; /* matched segments only; max(0, .) */
        }
        float fn = dur - ov;
        if (fn < 0.0f) fn = 0.0f;
        st->fn_sec += fn;
        st->p_sec += fn;
    }
    /* statistics.zig:104-111 */
    st->tpr = st->tp_sec / st->p_sec;
    st->fnr = st->fn_sec / st->p_sec;
    st->fdr = st->fp_sec / (st->fp_sec + st->tp_sec);
    st->ppv = st->tp_sec / (st->tp_sec + st->fp_sec);
    const float b2 = cfg->beta * cfg->beta;
    st->f_score = (1 + b2) * (st->ppv * st->tpr) / (b2 * st->ppv + st->tpr);
    st->fm = sqrtf(st->ppv * st->tpr);
    free(vad); free(ref); free(match);
}
