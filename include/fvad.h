/*
 * fvad.h -- C ABI of libfvad_cuda.so, the B200-native drop-in for Formula-VAD's per-frame
 * audio hot path (STFT -> NSNet2 -> ISTFT -> 1024-pt band energy -> VAD state machine).
 *
 * Every entry point names the reference interface it replaces (paths relative to the
 * reference checkout recursiveGecko/Formula-VAD @ 647d06f).  The reference is Zig; the
 * binding a maintainer would add is a set of `extern fn` declarations, shown in INTEGRATION.md.
 *
 * Conventions
 *   - plain pointers and sizes only; no exceptions, no aborts on bad arguments
 *   - every function that can fail returns int: FVAD_OK (0) or a negative fvad_status; the
 *     message of the last failure on the calling thread is fvad_last_error_string()
 *   - handles are thread-compatible: different handles may be used from different threads,
 *     one handle needs external synchronisation (the reference pipeline is single-threaded,
 *     src/simulator.zig:221-232 runs one OS thread per pipeline)
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     FVAD_ERR_NO_DEVICE
 */
#ifndef FVAD_H
#define FVAD_H

#include "fvad_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define FVAD_API
#else
#define FVAD_API __attribute__((visibility("default")))
#endif

/* One code per Zig error name on this path (SURVEY 8(b) "Errors"). */
typedef enum fvad_status {
    FVAD_OK = 0,
    FVAD_ERR_INVALID_SAMPLE_RATE = -1,    /* VADPipeline.zig:55-58  error.InvalidSampleRate */
    FVAD_ERR_INVALID_INPUT_LENGTH = -2,   /* NSNet2.zig:167-169     error.InvalidInputLength */
    FVAD_ERR_INVALID_FFT_SIZE = -3,       /* FFT.zig:41-43          error.InvalidFFTSize */
    FVAD_ERR_INVALID_SAMPLES_LENGTH = -4, /* FFT.zig:91-93          error.InvalidSamplesLength */
    FVAD_ERR_INVALID_WINDOW_LENGTH = -5,  /* FFT.zig:95-97          error.InvalidWindowLength */
    FVAD_ERR_INVALID_RESULT_LENGTH = -6,  /* FFT.zig:99-102,124-126 error.InvalidResultLength */
    FVAD_ERR_INVALID_BINS_LENGTH = -7,    /* FFT.zig:120-122        error.InvalidBinsLength */
    FVAD_ERR_OUT_OF_RANGE = -8,           /* FFT.zig:157-159,173-175 error.OutOfRange */
    FVAD_ERR_NEGATIVE_FREQUENCY = -9,     /* FFT.zig:161-163        error.NegativeFrequency */
    FVAD_ERR_CHANNEL_COUNT_MISMATCH = -10,/* SegmentWriter.zig:69-71 error.ChannelCountMismatch */
    FVAD_ERR_OUT_OF_MEMORY = -11,         /* error.OutOfMemory */
    FVAD_ERR_MODEL = -12,                 /* onnxruntime.zig:215-226 error.OnnxError (model load) */
    FVAD_ERR_CUDA = -13,                  /* a CUDA runtime call failed */
    FVAD_ERR_INVALID_ARGUMENT = -14,      /* NULL / misaligned pointer, bad index */
    FVAD_ERR_NO_DEVICE = -15              /* no CUDA device: this library has no CPU path */
} fvad_status;

FVAD_API const char *fvad_last_error_string(void);
FVAD_API const char *fvad_status_name(int status);
FVAD_API int fvad_device_count(int *count);

/* defaults of VADMachine.zig:30-51, VADPipeline.zig:20-26, AudioPipeline.zig:20-26 */
FVAD_API void fvad_vad_machine_config_default(fvad_vad_machine_config *cfg);
FVAD_API void fvad_config_default(fvad_config *cfg, size_t sample_rate, size_t n_channels);

/* ------------------------------------------------------------------------------------------
 * FFT operator -- replaces src/FFT.zig (kissfft wrapper).  GPU transform sizes: n_fft = 64*R
 * with R in {1,2,4,5,8,16} (64 ... 1024; the path uses 320 and 1024).
 * ---------------------------------------------------------------------------------------- */
typedef struct fvad_fft fvad_fft;

/* FFT.init, FFT.zig:35-75 */
FVAD_API int fvad_fft_create(size_t n_fft, size_t sample_rate, int mode_inverse, fvad_fft **out);
/* FFT.deinit, FFT.zig:77-82 */
FVAD_API void fvad_fft_destroy(fvad_fft *h);
/* FFT.fft, FFT.zig:85-113: `samples` is the two-piece SplitSlice (first, second), multiplied by
 * `window`, then an unnormalised real FFT into n_fft/2+1 bins.  Host pointers. */
FVAD_API int fvad_fft_forward(fvad_fft *h, const float *first, size_t n_first, const float *second,
                              size_t n_second, const float *window, size_t n_window,
                              fvad_complex *bins, size_t n_bins);
/* FFT.invFft, FFT.zig:115-134: unnormalised inverse (n_fft * true inverse, as kiss_fftri). */
FVAD_API int fvad_fft_inverse(fvad_fft *h, const fvad_complex *bins, size_t n_bins, float *result,
                              size_t n_result);
/* Batched forms (no reference counterpart; the reference loops, NSNet2.zig:248-263,326-338).
 * frames: n_frames frames, frame i starts at frames + i*frame_stride (overlap allowed);
 * window may be NULL (rectangular).  on_device: 0 = host pointers, 1 = device pointers. */
FVAD_API int fvad_fft_forward_batch(fvad_fft *h, const float *frames, size_t n_frames, size_t frame_stride,
                                    const float *window, fvad_complex *bins, int on_device);
FVAD_API int fvad_fft_inverse_batch(fvad_fft *h, const fvad_complex *bins, size_t n_frames, float *result,
                                    int on_device);
/* FFT.binCount / binWidth / nyquistFreq / freqToBin / binToFreq, FFT.zig:136-180 */
FVAD_API size_t fvad_fft_bin_count(const fvad_fft *h);
FVAD_API float fvad_fft_bin_width(const fvad_fft *h);
FVAD_API float fvad_fft_nyquist_freq(const fvad_fft *h);
FVAD_API int fvad_fft_freq_to_bin(const fvad_fft *h, float freq, size_t *bin);
FVAD_API int fvad_fft_bin_to_freq(const fvad_fft *h, size_t bin, float *freq);

/* ------------------------------------------------------------------------------------------
 * NSNet2 operator -- replaces src/NSNet2.zig (+ the onnxruntime session it owns).
 * ---------------------------------------------------------------------------------------- */
typedef struct fvad_nsnet2 fvad_nsnet2;

typedef enum fvad_nn_mode {
    FVAD_NN_F32 = 0,     /* fp32 CUDA-core GEMMs: the arithmetic of the reference (ORT CPU f32) */
    FVAD_NN_TC_F16X3 = 1,/* tcgen05, fp16 hi/lo split operands (3 MMAs), fp32 accumulate: ~fp32 accuracy */
    FVAD_NN_TC_F16 = 2   /* tcgen05, single fp16 MMA, fp32 accumulate: fast mode, own tolerance */
} fvad_nn_mode;

/* NSNet2.init, NSNet2.zig:35-141.  model_path NULL -> "data/nsnet2-20ms-baseline.onnx" (:56).
 * Accepted files: the FVADNS2 weight blob; .onnx is parsed by
 * tools/onnx_to_blob.py into that blob. */
FVAD_API int fvad_nsnet2_create(size_t sample_rate, const char *model_path, fvad_nsnet2 **out);
/* same, weights passed in memory (copied) */
FVAD_API int fvad_nsnet2_create_with_weights(size_t sample_rate, const fvad_nsnet2_weights *w, int nn_mode,
                                             fvad_nsnet2 **out);
/* NSNet2.deinit, NSNet2.zig:143-155 */
FVAD_API void fvad_nsnet2_destroy(fvad_nsnet2 *h);
/* NSNet2.getChunkSize, NSNet2.zig:157-159 */
FVAD_API size_t fvad_nsnet2_chunk_size(size_t sample_rate);
/* NSNet2.denoise, NSNet2.zig:161-237: one chunk of one channel; stateful (input/output halo,
 * 4 context feature frames, last_sample).  Host pointers. */
FVAD_API int fvad_nsnet2_denoise(fvad_nsnet2 *h, const float *first, size_t n_first, const float *second,
                                 size_t n_second, float *denoised_result, size_t n_result);
/* NSNet2.calcSpectrogram, NSNet2.zig:239-264: n_frames frames of 320 with hop 160 from
 * audio_chunk (len >= 160*n_frames+160), window len 320, result n_frames*161 bins.  Host pointers. */
FVAD_API int fvad_nsnet2_calc_spectrogram(const float *audio_chunk, size_t n_audio, size_t n_frames,
                                          const float *window, size_t n_window, fvad_complex *result,
                                          size_t n_result);

/* ------------------------------------------------------------------------------------------
 * Pipeline facade -- replaces src/AudioPipeline.zig (+ AudioPipeline/VADPipeline.zig and the
 * steps it drives).  Recording callbacks: for every VAD segment of the main machine the original and the
 * denoised PCM [sample_from, sample_to) are handed to the callbacks once those samples exist -- what
 * startRecording / endRecording and the two MRBRecorders deliver (AudioPipeline.zig:181-209,
 * AudioPipeline/MRBRecorder.zig:76-203).
 * ---------------------------------------------------------------------------------------- */
typedef struct fvad_pipeline fvad_pipeline;

/* AudioPipeline.Callbacks, AudioPipeline.zig:14-18 (recordings: planar f32, borrowed for the call) */
typedef struct fvad_recording {
    size_t n_channels, length, sample_rate;
    uint64_t global_start_frame_number;
    const float *const *channel_pcm;
} fvad_recording;
typedef struct fvad_callbacks {
    void *ctx;
    void (*on_original_recording)(void *ctx, const fvad_recording *rec);
    void (*on_denoised_recording)(void *ctx, const fvad_recording *rec);
} fvad_callbacks;

/* AudioPipeline.init, AudioPipeline.zig:40-102 (callbacks may be NULL) */
FVAD_API int fvad_pipeline_create(const fvad_config *cfg, const fvad_callbacks *callbacks, fvad_pipeline **out);
/* same with in-memory weights and an explicit NN mode */
FVAD_API int fvad_pipeline_create_with_weights(const fvad_config *cfg, const fvad_callbacks *callbacks,
                                               const fvad_nsnet2_weights *w, int nn_mode, fvad_pipeline **out);
/* AudioPipeline.deinit, AudioPipeline.zig:104-112 */
FVAD_API void fvad_pipeline_destroy(fvad_pipeline *p);
/* AudioPipeline.pushSamples, AudioPipeline.zig:118-143: planar PCM, n samples per channel;
 * returns the absolute index of the first pushed sample; runs every complete chunk. */
FVAD_API int fvad_pipeline_push_samples(fvad_pipeline *p, const float *const *channel_pcm, size_t n_channels,
                                        size_t n_samples, uint64_t *first_sample_index);
/* AudioPipeline.totalWriteCount, AudioPipeline.zig:114-116 */
FVAD_API uint64_t fvad_pipeline_total_write_count(const fvad_pipeline *p);
/* read-out of `pipeline.vad.vad_machine.vad_segments` (VADMachine.zig:73; used at
 * simulator/SimulationInstance.zig:221); machine 0 = main, 1.. = alt_vad_machines (VADPipeline.zig:24).
 * The array is owned by the pipeline and valid until the next push/destroy. */
FVAD_API int fvad_pipeline_segments(fvad_pipeline *p, size_t machine, const fvad_segment **segments, size_t *n);

/* ------------------------------------------------------------------------------------------
 * Batch engine -- many independent pipelines advanced in lock step on one GPU.  No reference
 * counterpart as a type: it is the simulator's "one thread per instance" (simulator.zig:221-232)
 * folded into the GPU batch dimension.  fvad_pipeline is a batch of one stream.
 * ---------------------------------------------------------------------------------------- */
typedef struct fvad_batch fvad_batch;

typedef struct fvad_batch_options {
    int32_t device;            /* CUDA device ordinal */
    int32_t nn_mode;           /* fvad_nn_mode */
    int32_t keep_taps;         /* 1: keep spectra/features/gains/denoised of the last push for fvad_batch_tap */
    int32_t stft_only;         /* 1: stop after STFT + features (BASELINE config 2) */
    size_t max_span_chunks;    /* chunks per stream processed per internal span (0 = auto) */
    const float *forced_gains; /* optional, host: [n_streams*n_channels][frames][161]; skips the network */
    size_t forced_gain_frames;
} fvad_batch_options;

typedef struct fvad_batch_stats {
    uint64_t kernel_launches;  /* kernels of this library launched since creation */
    uint64_t chunks_processed; /* stream-channel chunks */
    uint64_t vad_frames;       /* per stream */
    double last_push_device_ms;/* CUDA-event time of the last push, H2D/D2H included when host memory */
    double last_stft_ms, last_nn_ms, last_istft_ms, last_vadfft_ms, last_scan_ms; /* per-stage CUDA-event time, summed over the spans of the last push */
    /* NSNet2 per layer kernel {pack, fc1, W1, gru1, W2, gru2, fc2, fc3, fc4}: CUDA-event time and launches of the last push */
    double last_nn_layer_ms[9];
    uint32_t last_nn_layer_launches[9];
} fvad_batch_stats;

typedef enum fvad_tap {
    FVAD_TAP_SPECTRUM = 0, /* fvad_complex [sc][frames][161], before gain */
    FVAD_TAP_FEATURES = 1, /* float [sc][frames][161] */
    FVAD_TAP_GAINS = 2,    /* float [sc][frames][161] */
    FVAD_TAP_DENOISED16 = 3,/* float [sc][8000*chunks] denoised signal at 16 kHz */
    FVAD_TAP_BAND = 4,     /* float [streams][vad_frames] min-over-channels band sum (machine 0 band) */
    FVAD_TAP_RATIO = 5,    /* float [streams][vad_frames] sample-weighted volume ratio */
    FVAD_TAP_MET = 6,      /* uint8 [streams][vad_frames] threshold_met of machine 0 */
    FVAD_TAP_DENOISED = 7  /* float [sc][chunk*chunks] denoised signal at the input rate (1:D upsample) */
} fvad_tap;

FVAD_API void fvad_batch_options_default(fvad_batch_options *opt);
FVAD_API int fvad_batch_create(size_t n_streams, const fvad_config *cfg, const fvad_nsnet2_weights *w,
                               const fvad_batch_options *opt, fvad_batch **out);
FVAD_API void fvad_batch_destroy(fvad_batch *b);
/* pcm: n_streams*n_channels planar pointers (stream-major), n_samples each.  memory: 0 host,
 * 1 device (16-byte aligned).  Every complete chunk of every stream is processed.  Host memory is
 * ingested through two device stage buffers so the copy of span i+1 overlaps the kernels of span i
 * (pinned host memory makes the copy asynchronous). */
FVAD_API int fvad_batch_push(fvad_batch *b, const float *const *pcm, size_t n_pointers, size_t n_samples,
                             int memory);
/* same with 16-bit PCM, converted on the device as x / 32768 -- the normalisation libsndfile applies when the
 * reference reads its WAV input as float (src/audio_utils/AudioFileStream.zig:56-102,
 * simulator/SimulationInstance.zig:183-213); halves the host->device bytes. */
FVAD_API int fvad_batch_push_i16(fvad_batch *b, const int16_t *const *pcm, size_t n_pointers, size_t n_samples,
                                 int memory);
FVAD_API int fvad_batch_segments(fvad_batch *b, size_t stream, size_t machine, const fvad_segment **segments,
                                 size_t *n);
/* sizes are returned in elements of the tap's type for the LAST push; dst may be NULL to query */
FVAD_API int fvad_batch_tap(fvad_batch *b, int tap, void *dst, size_t dst_bytes, size_t *n_elements);
FVAD_API int fvad_batch_stats_get(fvad_batch *b, fvad_batch_stats *out);

/* ------------------------------------------------------------------------------------------
 * Evaluator -- replaces src/Evaluator.zig + src/Evaluator/statistics.zig (segment matching, TP/FP/FN
 * seconds, rates, F-score, Fowlkes-Mallows), f32 arithmetic in the reference's summation order.
 * ---------------------------------------------------------------------------------------- */
/* statistics.zig:77-83 defaults, beta = 0.7 */
FVAD_API void fvad_eval_config_default(fvad_eval_config *cfg);
/* Evaluator.initAndRun (Evaluator.zig:90-156) + statistics.fromEvaluator (statistics.zig:85-114) for one pair
 * of lists (host pointers, any order: both are sorted by start like the reference does). */
FVAD_API int fvad_evaluate(const fvad_eval_segment *vad, size_t n_vad, const fvad_eval_segment *reference, size_t n_reference,
                           const fvad_eval_config *cfg, fvad_eval_stats *out);
/* The same for every (stream, machine) of a batch in one launch, straight from the device-resident segment
 * lists (sample indices become seconds as f32(sample)/f32(sample_rate), SimulationInstance.zig:236-239).
 * references: all streams' labelled segments concatenated; reference_offsets[n_streams+1] delimits them.
 * out: [n_streams][n_machines].  This is the scoring half of the config sweep `alt_vad_machine_configs`
 * exists for (VADPipeline.zig:24). */
FVAD_API int fvad_batch_evaluate(fvad_batch *b, const fvad_eval_segment *references, const size_t *reference_offsets,
                                 const fvad_eval_config *cfg, fvad_eval_stats *out);

/* ------------------------------------------------------------------------------------------
 * VADMachine as an operator -- replaces VADMachine.init + VADMachine.run (src/AudioPipeline/VADMachine.zig:75-106,
 * 138-325) over a prepared series, for n_machines configurations at once (one warp each).  The inputs are what
 * run() derives from a VADPipeline.Result before it touches any state: band_min[m][f] is machine m's speech-band
 * sum minimised over the channels (:141-158), ratio[f] the frame's channel volume ratio (:173).  Frame f starts at
 * sample f * fft_size.  frames_per_launch > 0 cuts the series into launches of that many frames (the state is
 * carried in device memory in between, as the pipeline does from span to span); 0 = one launch.
 * met: [n_machines][n_frames] threshold decisions or NULL; segments: [n_machines][max_segments];
 * n_segments[m] counts every completed segment, also those beyond max_segments.  Host pointers.
 * ---------------------------------------------------------------------------------------- */
FVAD_API int fvad_vad_machine_scan(const fvad_vad_machine_config *cfgs, size_t n_machines, size_t sample_rate,
                                   size_t fft_size, const float *band_min, const float *ratio, size_t n_frames,
                                   size_t frames_per_launch, uint8_t *met, fvad_segment *segments, size_t max_segments,
                                   size_t *n_segments);

/* ------------------------------------------------------------------------------------------
 * fvad_vad_machine_scan without device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same kernels and results as fvad_vad_machine_scan, but:
 *   - every data pointer is a DEVICE pointer supplied by the caller: d_band_min [n_machines][n_frames], d_ratio [n_frames],
 *     d_met [n_machines][n_frames] or NULL, d_segments [n_machines][max_segments], d_n_segments [n_machines] (uint64:
 *     every completed segment, also those beyond max_segments); cfgs is a host array;
 *   - the operator's scratch (derived machine table, scan state, the three rolling-average rings per machine) is a
 *     file-scope static __device__ array (formula-vad_b200/csrc/noalloc.cu), ~2.1 MB that every GPU which loads the library
 *     holds until the process exits.  Limits: n_machines <= FVAD_NOALLOC_MAX_MACHINES and, per machine,
 *     long + short + ratio window lengths in frames <= FVAD_NOALLOC_MAX_RING (the defaults at 48 kHz need 8 469);
 *     beyond them the call fails with FVAD_ERR_INVALID_ARGUMENT and a message naming the limit;
 *   - the call allocates nothing on the device (no cudaMalloc / cudaFree, no allocating library routine): it launches
 *     three kernels on `cuda_stream` (a cudaStream_t, NULL = the default stream) of the current device and synchronises
 *     that stream before it returns.
 * The scratch is shared by every call and keeps the last call's contents; each call re-initialises it before the scan.
 * Calls that overlap on the GPU (two streams, two host threads) RACE on it: serialise them.
 * ---------------------------------------------------------------------------------------- */
#define FVAD_NOALLOC_MAX_MACHINES 16
#define FVAD_NOALLOC_MAX_RING 16384
FVAD_API int fvad_vad_machine_scan_noalloc(const fvad_vad_machine_config *cfgs, size_t n_machines, size_t sample_rate,
                                           size_t fft_size, const float *d_band_min, const float *d_ratio, size_t n_frames,
                                           uint8_t *d_met, fvad_segment *d_segments, size_t max_segments,
                                           uint64_t *d_n_segments, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* FVAD_H */
