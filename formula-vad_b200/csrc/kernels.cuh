// kernels.cuh -- launch wrappers of the hot-path kernels (definitions in the .cu files).
#pragma once
#include "common.cuh"

namespace fvad {

// Constant tables shared by all kernels, built once per device (tables.cu).
struct Tables {
    float2 *T320 = nullptr;    // exp(-2*pi*i*j/320), j<320
    float2 *T1024 = nullptr;   // exp(-2*pi*i*j/1024), j<1024
    float *win320 = nullptr;   // sqrt(symmetric Hann(320))           NSNet2.zig:384-396
    float *win320n = nullptr;  // win320[i] * (1/320)                  NSNet2.zig:323,335
    float *hann1024 = nullptr; // periodic Hann(fft_size)             BufferedFFT.zig:95-97
    float vad_norm = 0.f;      // windowNormFactor(hann)/fft_size     BufferedFFT.zig:99
    void build();
    void release();
};

// K1: framing + sqrt-Hann + rFFT-320 + log-power features (+ per-chunk sum of squares of the raw input).
struct StftArgs {
    const float *const *x;  // [SC] device pointers; the span starts at x[sc] + x_offset, D*8000*K samples
    size_t x_offset;
    const float *halo_in;   // [SC][160] decimated samples preceding the span
    float *halo_out;        // [SC][160] written by the span's last chunk
    fvad_complex *spec;     // [SC][50K][161] or nullptr
    float *feat;            // [SC][4+50K][161]; rows 0..3 are the context carried by the caller
    float *sumsq;           // [SC][1+K]; this kernel fills slots 1..K
    void *feat16;           // optional: the same feature rows as fp16, [SC][4+50K][176] (bins 161..175 zero): what the
                            // tensor-core backend reads; `feat` may then be nullptr (no f32 rows written)
                            // (core-matrix tiled, nn_tc.cuh); each frame is written as row f+4 of its own sequence and,
                            // for f >= 46, as context row f-46 of the next chunk's sequence (NSNet2.zig:188-203)
    int K, SC, D;
};
void launch_stft(const StftArgs &a, const Tables &t, cudaStream_t s);

// K3a: gain clamp * spectrum -> irFFT-320 -> window/320 -> overlap-add -> y16
struct IstftArgs {
    const fvad_complex *spec;  // [SC][50K][161]
    const float *gains;        // [SC][50K][161]
    const float *tail_in;      // [SC][160] second half of the frame preceding the span (windowed, scaled)
    float *tail_out;           // [SC][160]
    float *y16;                // [SC][kYLook + 8000K]; the span's samples start at +kYLook
    int K, SC;
};
void launch_istft(const IstftArgs &a, const Tables &t, cudaStream_t s);

// K3b: 1:D linear upsample -> periodic Hann(1024) -> rFFT-1024 -> |X|/512 -> band sums -> min over channels
struct VadFftArgs {
    const float *y16;     // [S*C][kYLook + 8000K]
    const float *sumsq;   // [S*C][1+K]  slot 0 = last chunk of the previous span
    float *band;          // [n_bands][S][max_frames]
    float *ratio;         // [S][max_frames]
    const int *band_lo, *band_hi;  // [n_bands] device
    int n_bands;
    int S, C, K, D;
    long long span_start;  // absolute sample index (input rate) of the span start
    long long first_frame; // absolute index of the first VAD frame to produce
    int n_frames;          // frames to produce
    int max_frames;        // row pitch of band/ratio
    int fft_size;          // 1024
    int chunk;             // D*8000
    int has_prev_chunk;    // slot 0 of sumsq valid
    unsigned long long *work_ctr = nullptr;  // device scratch for the dynamic item queue (nullptr: static striding)
};
void launch_vadfft(const VadFftArgs &a, const Tables &t, cudaStream_t s);

// K4: VAD state machine scan, one thread per (stream, machine)
struct ScanMachine {  // derived per machine on the host from fvad_vad_machine_config (VADMachine.zig:75-106,160-163)
    int band;                  // index into band rows
    int n_long, n_short, n_ratio;
    int has_init; double init_avg;
    double threshold_factor, ratio_threshold;
    unsigned long long min_to_open, max_gap, start_margin, end_margin;
    float min_duration_sec, frame_sec, sample_rate_f;
};
struct ScanState {  // carried across pushes, one per (stream, machine)
    int state;
    int long_idx, long_cnt, short_idx, short_cnt, ratio_idx, ratio_cnt;
    int has_last;
    double long_sum, short_sum, ratio_sum, last_avg;
    unsigned long long start_index, end_index;
    float vol_ratio_sum; unsigned long long vol_ratio_count; float met_sec;
    unsigned int n_segments;
};
struct ScanArgs {
    const float *band, *ratio;   // see VadFftArgs
    const ScanMachine *machines; // [M]
    ScanState *states;           // [S][M]
    double *rings;               // [S][M][ring_pitch]: long | short | ratio
    int ring_pitch;
    fvad_segment *segments;      // [S][M][max_segments]
    int max_segments;
    unsigned char *met;          // [S][max_frames] machine-0 decisions or nullptr
    int met_per_machine = 0;     // 1: met is [S][M][max_frames], every machine writes its row
    int S, M, n_frames, max_frames;
    long long first_frame;       // absolute frame number of column 0
    int fft_size;
    int fast;                    // every machine has n_long >= 514 and n_short + n_ratio <= 64 (vadscan.cu)
};
void launch_vadscan(const ScanArgs &a, cudaStream_t s);

// small helpers
void launch_copy_rows(float *dst, const float *src, size_t dst_pitch, size_t src_pitch, size_t row_len, size_t rows, cudaStream_t s);
void launch_i16_to_f32(float *dst, const int16_t *src, size_t dst_pitch, size_t src_pitch, size_t n, size_t rows, cudaStream_t s);

}  // namespace fvad
