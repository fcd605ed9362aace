// engine.cuh -- the batch engine: S independent pipelines (streams) x C channels advanced in lock
// step on one GPU.  Host-side restatement of the step order of the reference's
// AudioPipeline.pushSamples (src/AudioPipeline.zig:118-143) -> VADPipeline.collectInputStep /
// denoiserStep / fftStep / stateMachineStep (src/AudioPipeline/VADPipeline.zig:140-237), with time
// folded into the batch: every complete chunk of every stream-channel is processed in one pass
// K1 (stft) -> NN -> K3a (istft) -> K3b (vadfft) -> K4 (vadscan); the only state carried between
// passes is the halo set of SURVEY App. B (160 input samples, 4 feature rows, 160 output samples,
// 1024 denoised samples, one chunk ratio, the VAD machine states).
#pragma once
#include <algorithm>
#include <array>
#include <memory>
#include <vector>

#include "kernels.cuh"
#include "nn.cuh"

namespace fvad {

struct MachineDerived {
    ScanMachine dev;
    int bin_lo, bin_hi;
};

class Batch {
public:
    Batch(size_t n_streams, const fvad_config &cfg, const HostWeights *weights, const fvad_batch_options &opt);
    ~Batch();

    // pcm: S*C planar pointers (stream-major), n samples each; memory 0 host / 1 device
    void push(const float *const *pcm, size_t n_pointers, size_t n_samples, int memory) { push_any(reinterpret_cast<const void *const *>(pcm), n_pointers, n_samples, memory, 0); }
    // same with 16-bit PCM (x / 32768, what libsndfile's sf_readf_float hands the reference: AudioFileStream.zig:56-102)
    void push_i16(const int16_t *const *pcm, size_t n_pointers, size_t n_samples, int memory) { push_any(reinterpret_cast<const void *const *>(pcm), n_pointers, n_samples, memory, 1); }
    void segments(size_t stream, size_t machine, const fvad_segment **segs, size_t *n);
    size_t tap(int which, void *dst, size_t dst_bytes);
    void stats(fvad_batch_stats *out) const;
    // eval.cu: statistics of every (stream, machine) list against the streams' reference segments; out [S][M]
    void evaluate(const fvad_eval_segment *refs, const size_t *ref_offsets, const fvad_eval_config &cfg, fvad_eval_stats *out);
    uint64_t total_write_count() const { return total_written_; }
    bool skipping() const { return skip_; }
    uint64_t processed_samples() const { return processed_; }   // per channel, whole chunks
    // FSM state (0 closed, 1 opening, 2 open, 3 closing) and speech_start_index of a machine after the last push
    void machine_state(size_t stream, size_t machine, int *state, uint64_t *start_index) const {
        const ScanState &st = h_states_[stream * (size_t)M_ + machine];
        if (state) *state = st.state;
        if (start_index) *start_index = st.start_index;
    }
    size_t n_machines() const { return (size_t)M_; }
    size_t n_streams() const { return (size_t)S_; }
    size_t n_channels() const { return (size_t)C_; }
    size_t chunk() const { return (size_t)chunk_; }

private:
    void push_any(const void *const *pcm, size_t n_pointers, size_t n_samples, int memory, int fmt);
    void process_span(const float *const *d_x, size_t x_offset, int K);
    void init_machines(const fvad_config &cfg);
    void fetch_segments();

    int S_, C_, SC_, D_, chunk_, M_, fft_size_;
    size_t sample_rate_;
    fvad_batch_options opt_;
    int Kmax_;
    cudaStream_t stream_ = nullptr, scan_stream_ = nullptr, copy_stream_ = nullptr;
    cudaEvent_t scan_done_ = nullptr;          // the most recently queued scan
    cudaEvent_t scan_done_b_[2] = {};          // the scan that last read band_/ratio_/met_[b]
    cudaEvent_t after_rnn_ = nullptr;          // NnSpan::after_rnn of the span in flight
    // The scan of span i is queued behind the recurrent layers of span i+1 (see process_span), so it is held here
    struct PendingScan { bool valid = false; int buf = 0, n_frames = 0; long long first_frame = 0; size_t span = 0; } pscan_;
    int vbuf_ = 0;                              // band_/ratio_/met_ buffer the next span's K3b writes
    bool defer_scan_ = true;
    void issue_scan(cudaEvent_t also_after);
    cudaEvent_t stage_ready_[2] = {}, stage_free_[2] = {};  // double-buffered ingest: copy stream <-> compute stream
    cudaEvent_t ev_[8] = {};
    std::vector<std::array<cudaEvent_t, 7>> span_ev_;  // per-span stage boundaries of the current push
    size_t span_idx_ = 0;
    Tables tables_;
    std::unique_ptr<NnBackend> nn_;
    std::vector<MachineDerived> machines_;
    int n_bands_ = 0;
    int scan_fast_ = 1;

    // progress
    uint64_t total_written_ = 0;   // samples pushed per channel
    uint64_t processed_ = 0;       // samples consumed in whole chunks (per channel)
    uint64_t frames_done_ = 0;     // VAD frames produced per stream
    int n_pend_ = 0;               // samples waiting in pend_ (per channel)
    int parity_ = 0;               // which halo/tail carry buffer is current

    // device state
    DevBuf<float> pend_, stage_[2];            // [SC][chunk], 2 x [SC][Kstage*chunk]
    DevBuf<int16_t> pend_i16_, stage_i16_[2];             // raw 16-bit PCM of a span before conversion
    DevBuf<const float *> d_xptr_, d_sptr_[2]; // [SC] caller's device pointers / stage buffer rows
    size_t stage_pitch_ = 0;
    DevBuf<float> halo_[2], tail_[2];          // [SC][160]
    DevBuf<float> ctx_carry_, ylook_carry_, sumsq_carry_;
    DevBuf<float> feat16_, ctx16_carry_;  // fp16 feature rows [SC][4+50K][176] and their 4-row carry, held as 88 floats per row
    DevBuf<fvad_complex> spec_;
    DevBuf<float> feat_, gains_, y16_, sumsq_, band_[2], ratio_[2];
    DevBuf<unsigned char> met_[2];
    DevBuf<unsigned long long> work_ctr_;  // dynamic work queues of the persistent spectral kernels
    DevBuf<int> band_lo_, band_hi_;
    DevBuf<ScanMachine> d_machines_;
    DevBuf<ScanState> d_states_;
    DevBuf<double> rings_;
    DevBuf<fvad_segment> d_segments_;
    DevBuf<float> forced_;
    int ring_pitch_ = 0, max_segments_ = 0, max_frames_ = 0;

    // host results
    std::vector<ScanState> h_states_;
    std::vector<fvad_segment> h_segments_;
    std::vector<unsigned> seg_seen_;
    bool segment_overflow_ = false;
    bool skip_ = false;

    // taps of the last push
    struct Taps {
        size_t frames = 0, chunks = 0, vad_frames = 0;  // per stream-channel / per stream
        DevBuf<fvad_complex> spec;
        DevBuf<float> feat, gains, y16, den, band, ratio;
        DevBuf<unsigned char> met;
        size_t fill_chunks = 0, fill_vad = 0, cap_chunks = 0, cap_vad = 0;
    } taps_;

    // stats
    uint64_t launches0_ = 0, chunks_processed_ = 0;
    double last_ms_[6] = {};
    double last_layer_ms_[9] = {};
    unsigned last_layer_launches_[9] = {};
};

// eval.cu: one pair of lists (host pointers)
ScanMachine derive_machine(const fvad_vad_machine_config &c, size_t sample_rate, size_t fft_size);
void vad_machine_scan(const fvad_vad_machine_config *cfgs, size_t n_machines, size_t sample_rate, size_t fft_size,
                      const float *band_min, const float *ratio, size_t n_frames, size_t frames_per_launch,
                      unsigned char *met, fvad_segment *segments, size_t max_segments, size_t *n_segments);
void evaluate_lists(const fvad_eval_segment *vad, size_t n_vad, const fvad_eval_segment *ref, size_t n_ref,
                    const fvad_eval_config &cfg, fvad_eval_stats *out);

}  // namespace fvad
