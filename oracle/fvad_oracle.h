/*
 * fvad_oracle.h -- CPU ORACLE for the Formula-VAD per-frame audio hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library; the product library
 * (libfvad_cuda.so) never links, loads or calls anything in oracle/.
 *
 * Every function is a plain-C restatement of a reference function; the reference
 * file:line is cited at each definition (paths relative to the reference checkout).
 *
 * Parity status:
 *   - FFT (rows a6/a10/a14): PINNED against oracle/_ref/libkissfft_ref.so, the reference's own
 *     kissfft C compiled unmodified (tests/test_oracle_fft.py).
 *   - VADMetadata / statistics: PINNED against the reference's own unit-test KATs
 *     (VADMetadata.zig:70-110, statistics.zig:286-360).
 *   - NSNet2 inference (row a8): PARITY UNPINNED -- onnxruntime 1.15 and the .onnx weights are
 *     absent from the reference snapshot; the restatement is cross-checked against torch.nn.GRU only.
 *   - VADMachine / RollingAverage / DSP glue: no reference golden vectors exist;
 *     restated line by line, unpinned beyond that.
 */
#ifndef FVAD_ORACLE_H
#define FVAD_ORACLE_H

#include "../include/fvad_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- FFT (oracle_fft.c) ---- */
typedef struct orc_fft orc_fft;
typedef struct orc_rfft orc_rfft;
orc_fft *orc_fft_new(int n, int inverse);
void orc_fft_free(orc_fft *p);
void orc_fft_run(const orc_fft *p, const fvad_complex *in, fvad_complex *out);
orc_rfft *orc_rfft_new(int n, int inverse);
void orc_rfft_free(orc_rfft *p);
int orc_rfft_size(const orc_rfft *p);
void orc_rfft_forward(orc_rfft *p, const float *x, fvad_complex *X);
void orc_rfft_inverse(orc_rfft *p, const fvad_complex *X, float *x);

/* Optional: route the real FFTs through the reference's own kissfft (oracle/_ref/libkissfft_ref.so).
 * Pass the three entry points obtained with dlsym; NULL restores the restatement. */
typedef void *(*orc_kiss_alloc_fn)(int nfft, int inverse, void *mem, size_t *lenmem);
typedef void (*orc_kiss_fftr_fn)(void *cfg, const float *timedata, fvad_complex *freqdata);
typedef void (*orc_kiss_fftri_fn)(void *cfg, const fvad_complex *freqdata, float *timedata);
void orc_use_kissfft(orc_kiss_alloc_fn alloc, orc_kiss_fftr_fn fwd, orc_kiss_fftri_fn inv);

/* ---- DSP pieces (oracle_dsp.c) ---- */
void orc_hann_periodic(float *w, int n);      /* window_fn.zig:22-28,51-68 */
void orc_hann_symmetric(float *w, int n);     /* window_fn.zig:30-41 */
float orc_window_norm_factor(const float *w, int n); /* window_fn.zig:8-16 */
void orc_nsnet2_window(float *w);             /* NSNet2.zig:384-396, 320 taps */
float orc_rms_volume(const float *x, size_t n); /* audio_utils.zig:14-24 */
void orc_downsample(const float *in, size_t n_in, float *out, int rate); /* resample.zig:9-29 */
float orc_upsample(const float *in, size_t n_in, float *out, float prev_last, int rate); /* resample.zig:32-79 */
int orc_freq_to_bin(float freq, size_t sample_rate, size_t n_fft); /* FFT.zig:156-167; <0 on error */

/* windowed forward real FFT as FFT.zig:85-113 does it (window multiply in f32, then kiss_fftr) */
typedef struct orc_fftop orc_fftop;
orc_fftop *orc_fftop_new(int n_fft, int inverse);
void orc_fftop_free(orc_fftop *p);
void orc_fftop_forward(orc_fftop *p, const float *samples, const float *window, fvad_complex *bins);
void orc_fftop_inverse(orc_fftop *p, const fvad_complex *bins, float *result);

/* ---- NSNet2 network (oracle_nn.c) ---- */
typedef struct orc_nn orc_nn;
orc_nn *orc_nn_new(const fvad_nsnet2_weights *w); /* copies the weights */
void orc_nn_free(orc_nn *p);
/* features [steps][161] -> gains [steps][161], zero initial GRU state (NSNet2.zig:220) */
void orc_nn_forward(orc_nn *p, const float *features, float *gains, int steps);
/* same network evaluated in double precision, for tolerance studies */
void orc_nn_forward_f64(const fvad_nsnet2_weights *w, const float *features, double *gains, int steps);

/* ---- NSNet2 denoiser object (oracle_dsp.c), one per channel: NSNet2.zig:35-237 ---- */
typedef struct orc_nsnet2 orc_nsnet2;
orc_nsnet2 *orc_nsnet2_new(size_t sample_rate, const fvad_nsnet2_weights *w);
void orc_nsnet2_free(orc_nsnet2 *p);
size_t orc_nsnet2_chunk_size(size_t sample_rate); /* NSNet2.zig:157-159 */
int orc_nsnet2_denoise(orc_nsnet2 *p, const float *samples, size_t n, float *out);
/* taps of the last denoise() call, for parity tests */
const fvad_complex *orc_nsnet2_last_spec(const orc_nsnet2 *p); /* [50][161], BEFORE gain */
const float *orc_nsnet2_last_features(const orc_nsnet2 *p);    /* [54][161] */
const float *orc_nsnet2_last_gains(const orc_nsnet2 *p);       /* [54][161] */
/* if set, denoise() skips the network and uses these gains [50][161] for the new frames */
void orc_nsnet2_force_gains(orc_nsnet2 *p, const float *gains50);

/* ---- VAD (oracle_vad.c) ---- */
typedef struct orc_rolling orc_rolling;                    /* structures/RollingAverage.zig */
orc_rolling *orc_rolling_new(size_t count, int has_init, double init);
void orc_rolling_free(orc_rolling *p);
double orc_rolling_push(orc_rolling *p, float v);
int orc_rolling_last(const orc_rolling *p, double *out); /* 0 if last_avg is null */

typedef struct orc_meta { /* AudioPipeline/VADMetadata.zig: optional fields as flag+value */
    int has_ratio; float ratio_sum, ratio_weight;
    int has_min; float vmin;
    int has_max; float vmax;
} orc_meta;
typedef struct orc_meta_result {
    int has_ratio; float ratio;
    int has_min; float vmin;
    int has_max; float vmax;
} orc_meta_result;
void orc_meta_push(orc_meta *m, const orc_meta_result *values, float weight); /* VADMetadata.zig:29-60 */
orc_meta_result orc_meta_to_result(const orc_meta *m);                         /* VADMetadata.zig:16-27 */

typedef struct orc_vadm orc_vadm; /* AudioPipeline/VADMachine.zig */
orc_vadm *orc_vadm_new(const fvad_vad_machine_config *cfg, size_t sample_rate, size_t n_channels, size_t fft_size);
void orc_vadm_free(orc_vadm *p);
/* one VAD frame: per-channel band sums, frame start index, metadata ratio (has_ratio=0 -> null).
 * returns recording_state (0 none,1 started,2 completed,3 aborted); *sample_number as the reference. */
int orc_vadm_run(orc_vadm *p, const float *channel_band, uint64_t index, int has_ratio, float ratio, uint64_t *sample_number);
size_t orc_vadm_segments(const orc_vadm *p, const fvad_segment **segs);
int orc_vadm_last_met(const orc_vadm *p);   /* threshold_met of the last run() */
int orc_vadm_state(const orc_vadm *p);      /* 0 closed 1 opening 2 open 3 closing */

/* ---- whole pipeline (oracle_pipeline.c): AudioPipeline.pushSamples -> VADPipeline.run ---- */
typedef struct orc_pipeline orc_pipeline;
/* allow_any_rate: 0 = reject sample_rate != 48000 like VADPipeline.zig:55-58;
 *                 1 = documented deviation, also accept 16000. */
orc_pipeline *orc_pipeline_new(const fvad_config *cfg, const fvad_nsnet2_weights *w, int allow_any_rate, int *err);
void orc_pipeline_free(orc_pipeline *p);
int orc_pipeline_push(orc_pipeline *p, const float *const *channel_pcm, size_t n, uint64_t *first_index);
uint64_t orc_pipeline_total_write_count(const orc_pipeline *p);
size_t orc_pipeline_segments(const orc_pipeline *p, size_t machine, const fvad_segment **segs);
/* per-VAD-frame taps (grow as frames are produced): band min over channels, metadata ratio,
 * threshold_met of machine 0 */
size_t orc_pipeline_frames(const orc_pipeline *p, const float **band_min, const float **ratio, const uint8_t **met);
/* denoised output tap: keeps the last `keep` chunks per channel if enabled (0 = off) */
void orc_pipeline_keep_denoised(orc_pipeline *p, int on);
size_t orc_pipeline_denoised(const orc_pipeline *p, size_t channel, const float **pcm);
/* feature/gain/spectrum taps of channel 0 accumulated over all chunks when enabled */
void orc_pipeline_keep_nn_taps(orc_pipeline *p, int on);
size_t orc_pipeline_nn_taps(const orc_pipeline *p, const float **features50, const float **gains50, const fvad_complex **spec50);
/* if set, every denoiser uses these gains instead of running the network: [n_channels][n_frames][161] */
void orc_pipeline_force_gains(orc_pipeline *p, const float *gains, size_t n_frames_per_channel);

/* ---- Evaluator (oracle_eval.c): Evaluator.zig:90-156 + Evaluator/statistics.zig ---- */
typedef struct orc_eval_seg { float from_sec, to_sec; } orc_eval_seg;
typedef struct orc_eval_cfg { float extrude_start, extrude_end, fill_gaps, ignore_shorter_than, beta; } orc_eval_cfg;
typedef struct orc_eval_stats {
    float p_sec, tp_sec, fp_sec, fn_sec;
    float tpr, ppv, fnr, fdr, f_score, fm;
} orc_eval_stats;
void orc_evaluate(const orc_eval_seg *vad, size_t n_vad, const orc_eval_seg *ref, size_t n_ref,
                  const orc_eval_cfg *cfg, orc_eval_stats *out);
float orc_eval_false_positive_sec(const orc_eval_seg *vad_seg, const orc_eval_seg *refs, size_t n_refs,
                                  const orc_eval_cfg *cfg);

#ifdef __cplusplus
}
#endif
#endif
