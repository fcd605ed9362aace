/*
 * fvad_types.h -- plain-old-data types shared by the product C-ABI (fvad.h) and the
 * CPU oracle (oracle/, test infrastructure only).  No functions here.
 *
 * Every struct mirrors a reference type; the reference location is cited per field group.
 * Paths are relative to the reference checkout (recursiveGecko/Formula-VAD @ 647d06f).
 */
#ifndef FVAD_TYPES_H
#define FVAD_TYPES_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* src/FFT.zig:12-15  `Complex = extern struct { r: f32, i: f32 }` (bit-compatible with kiss_fft_cpx / float2) */
typedef struct fvad_complex {
    float r;
    float i;
} fvad_complex;

/* src/AudioPipeline/VADPipeline.zig:28-33  `SpeechSegment` */
typedef struct fvad_segment {
    uint64_t sample_from;
    uint64_t sample_to;
    float avg_channel_vol_ratio;
    float vad_met_sec;
} fvad_segment;

/* src/AudioPipeline/VADMachine.zig:30-51  `Config` (same field names, same defaults via
 * fvad_vad_machine_config_default()).  `initial_long_term_avg: ?f64` becomes a flag + value. */
typedef struct fvad_vad_machine_config {
    float speech_min_freq;             /* 500  */
    float speech_max_freq;             /* 2000 */
    float long_term_speech_avg_sec;    /* 180  */
    int32_t has_initial_long_term_avg; /* 1    */
    double initial_long_term_avg;      /* 0.005 */
    float short_term_speech_avg_sec;   /* 0.2  */
    float speech_threshold_factor;     /* 10   */
    float channel_vol_ratio_avg_sec;   /* 0.5  */
    float channel_vol_ratio_threshold; /* 0.5  */
    float min_consecutive_sec_to_open; /* 0.2  */
    float max_speech_gap_sec;          /* 2    */
    float min_vad_duration_sec;        /* 0.7  */
} fvad_vad_machine_config;

/* src/AudioPipeline/VADPipeline.zig:20-26  `Config` */
typedef struct fvad_vad_config {
    size_t fft_size; /* 1024 */
    fvad_vad_machine_config vad_machine_config;
    const fvad_vad_machine_config *alt_vad_machine_configs; /* may be NULL */
    size_t n_alt_vad_machine_configs;
    const char *denoiser_model_path; /* NULL -> "data/nsnet2-20ms-baseline.onnx" (src/NSNet2.zig:56) */
} fvad_vad_config;

/* src/AudioPipeline.zig:20-26  `Config`.  buffer_length 0 means "null" -> sample_rate * 10. */
typedef struct fvad_config {
    size_t sample_rate;
    size_t n_channels;
    size_t buffer_length;
    fvad_vad_config vad_config;
    int32_t skip_processing;
} fvad_config;

/* src/Evaluator/SpeechSegment.zig:7-9  `from_sec`, `to_sec` (the other fields are links and report text) */
typedef struct fvad_eval_segment {
    float from_sec;
    float to_sec;
} fvad_eval_segment;

/* src/Evaluator/statistics.zig:77-83  `StatConfig` (+ the F-score beta the reference fixes at 0.7, :108) */
typedef struct fvad_eval_config {
    float ignore_shorter_than_sec; /* 0 */
    float extrude_start;           /* 0 */
    float extrude_end;             /* 0 */
    float fill_gaps;               /* 0 */
    float f_score_beta;            /* 0.7 */
} fvad_eval_config;

/* src/Evaluator/statistics.zig:6-41  `SingleStats` (same field names) */
typedef struct fvad_eval_stats {
    float total_positives_sec;
    float true_positives_sec;
    float false_positives_sec;
    float false_negatives_sec;
    float true_positive_rate;
    float precision;
    float false_negative_rate;
    float false_discovery_rate;
    float f_score;
    float fm_index;
} fvad_eval_stats;

/* NSNet2 constants, src/NSNet2.zig:12-16 */
enum {
    FVAD_NS_NFFT = 320,
    FVAD_NS_HOP = 160,
    FVAD_NS_BINS = 161,
    FVAD_NS_FRAMES = 50,      /* chunk_size / n_hop */
    FVAD_NS_CTX = 4,          /* artifact_mitigation_window */
    FVAD_NS_STEPS = 54,       /* frames + ctx, the ORT tensor is [1,54,161] */
    FVAD_NS_CHUNK16 = 8000    /* chunk_size at 16 kHz */
};

/* NSNet2 baseline weights.  All row-major f32, ONNX layouts:
 *   fc  : w[out][in], b[out]
 *   gru : W[3H][in], R[3H][H], Wb[3H], Rb[3H]; gate order z, r, h; linear_before_reset = 1. */
typedef struct fvad_nsnet2_weights {
    int32_t n_bins;   /* 161 */
    int32_t hidden;   /* 400 */
    int32_t fc2;      /* 600 */
    int32_t fc3;      /* 600 */
    int32_t fc1_relu; /* 0: fc1 has no activation (to be confirmed against the .onnx node list) */
    const float *fc1_w, *fc1_b;
    const float *gru1_W, *gru1_R, *gru1_Wb, *gru1_Rb;
    const float *gru2_W, *gru2_R, *gru2_Wb, *gru2_Rb;
    const float *fc2_w, *fc2_b;
    const float *fc3_w, *fc3_b;
    const float *fc4_w, *fc4_b;
} fvad_nsnet2_weights;

#ifdef __cplusplus
}
#endif
#endif /* FVAD_TYPES_H */
