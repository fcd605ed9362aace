// capi.cu -- the extern "C" surface declared in include/fvad.h.  Exceptions never cross it.
#include <cmath>
#include <algorithm>
#include <cstring>
#include <new>

#include "engine.cuh"
#include "ops.cuh"

using namespace fvad;

namespace {

template <typename F>
int guarded(F &&f) {
    try {
        f();
        return FVAD_OK;
    } catch (const Error &e) {
        set_last_error(e.what());
        return e.code;
    } catch (const std::bad_alloc &) {
        set_last_error("out of host memory");
        return FVAD_ERR_OUT_OF_MEMORY;
    } catch (const std::exception &e) {
        set_last_error(e.what());
        return FVAD_ERR_INVALID_ARGUMENT;
    }
}

void require_device() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        cudaGetLastError();
        throw Error(FVAD_ERR_NO_DEVICE, "no CUDA device: libfvad_cuda has no CPU path");
    }
}

}  // namespace

struct fvad_fft {
    size_t n_fft, sample_rate;
    int inverse;
    FftPlan *plan = nullptr;
    DevBuf<float> d_real, d_win;
    DevBuf<fvad_complex> d_bins;
};
struct fvad_batch {
    Batch *impl = nullptr;
};
// Recording extraction for the pipeline facade: what AudioPipeline.startRecording / endRecording and the two
// MRBRecorders deliver (reference src/AudioPipeline.zig:181-209, AudioPipeline/MRBRecorder.zig:76-203): for every
// VAD segment of the main machine, the original and the denoised PCM [sample_from, sample_to) as planar f32,
// handed to the callbacks once the samples exist (sample_to lies 2 s past the speech end, so usually a later push
// completes it).  The reference copies out of its 10 s rings while a recording is in progress; here the facade keeps
// a host-side history that is trimmed to what a current or future recording can still reach.
struct RecHistory {
    std::vector<std::vector<float>> ch;
    uint64_t base = 0;  // absolute sample index of ch[c][0]
    uint64_t end() const { return base + (ch.empty() ? 0 : ch[0].size()); }
    void append(size_t c, const float *src, size_t n) { ch[c].insert(ch[c].end(), src, src + n); }
    void trim(uint64_t keep_from) {
        if (keep_from <= base) return;
        const size_t drop = (size_t)std::min<uint64_t>(keep_from - base, ch.empty() ? 0 : ch[0].size());
        for (auto &v : ch) v.erase(v.begin(), v.begin() + (long)drop);
        base += drop;
    }
};
struct PendingRec {
    uint64_t from, to;
    bool orig_done, den_done;
};
struct fvad_pipeline {
    Batch *impl = nullptr;
    fvad_callbacks cb{};
    bool has_cb = false;
    size_t sample_rate = 0, n_channels = 0, seg_seen = 0;
    RecHistory orig, den;
    std::vector<PendingRec> pending;
    std::vector<float> tapbuf;
};

static void deliver(const fvad_pipeline *p, const RecHistory &h, const PendingRec &r, bool denoised) {
    auto fn = denoised ? p->cb.on_denoised_recording : p->cb.on_original_recording;
    if (!fn) return;
    const uint64_t from = std::max(r.from, h.base);
    std::vector<const float *> ptrs(p->n_channels);
    for (size_t c = 0; c < p->n_channels; ++c) ptrs[c] = h.ch[c].data() + (from - h.base);
    fvad_recording rec;
    rec.n_channels = p->n_channels;
    rec.length = (size_t)(r.to - from);
    rec.sample_rate = p->sample_rate;
    rec.global_start_frame_number = from;
    rec.channel_pcm = ptrs.data();
    fn(p->cb.ctx, &rec);  // borrowed for the call, like the reference's AudioBuffer (AudioPipeline.zig:193-199)
}

static void run_recorders(fvad_pipeline *p, const float *const *pcm, size_t n) {
    const size_t C = p->n_channels;
    for (size_t c = 0; c < C; ++c) p->orig.append(c, pcm[c], n);
    // denoised samples of the chunks this push completed
    const size_t n_el = p->impl->tap(FVAD_TAP_DENOISED, nullptr, 0);
    if (n_el) {
        p->tapbuf.resize(n_el);
        p->impl->tap(FVAD_TAP_DENOISED, p->tapbuf.data(), n_el * sizeof(float));
        const size_t per = n_el / C;
        for (size_t c = 0; c < C; ++c) p->den.append(c, p->tapbuf.data() + c * per, per);
    }
    const fvad_segment *segs = nullptr;
    size_t n_seg = 0;
    p->impl->segments(0, 0, &segs, &n_seg);
    for (; p->seg_seen < n_seg; ++p->seg_seen)
        p->pending.push_back(PendingRec{segs[p->seg_seen].sample_from, segs[p->seg_seen].sample_to, false, false});
    for (auto &r : p->pending) {
        if (!r.orig_done && r.to <= p->orig.end()) { deliver(p, p->orig, r, false); r.orig_done = true; }
        if (!r.den_done && r.to <= p->den.end()) { deliver(p, p->den, r, true); r.den_done = true; }
    }
    p->pending.erase(std::remove_if(p->pending.begin(), p->pending.end(), [](const PendingRec &r) { return r.orig_done && r.den_done; }),
                     p->pending.end());
    // retention: unfinished recordings, a speech run in progress (its recording starts 2 s before it), and the 2 s
    // a run that starts in not-yet-processed audio can reach back
    const uint64_t margin = (uint64_t)((float)p->sample_rate * 2);
    uint64_t keep = p->impl->processed_samples();
    keep = keep > margin + 1024 ? keep - margin - 1024 : 0;
    int state = 0;
    uint64_t start = 0;
    p->impl->machine_state(0, 0, &state, &start);
    if (state != 0) keep = std::min(keep, start > margin ? start - margin : 0);
    for (const auto &r : p->pending) keep = std::min(keep, r.from);
    p->orig.trim(keep);
    p->den.trim(keep);
}
struct fvad_nsnet2 {
    Batch *impl = nullptr;
    size_t sample_rate = 0;
    std::vector<float> joined;
};

extern "C" {

const char *fvad_last_error_string(void) { return last_error().c_str(); }

const char *fvad_status_name(int s) {
    switch (s) {
        case FVAD_OK: return "OK";
        case FVAD_ERR_INVALID_SAMPLE_RATE: return "InvalidSampleRate";
        case FVAD_ERR_INVALID_INPUT_LENGTH: return "InvalidInputLength";
        case FVAD_ERR_INVALID_FFT_SIZE: return "InvalidFFTSize";
        case FVAD_ERR_INVALID_SAMPLES_LENGTH: return "InvalidSamplesLength";
        case FVAD_ERR_INVALID_WINDOW_LENGTH: return "InvalidWindowLength";
        case FVAD_ERR_INVALID_RESULT_LENGTH: return "InvalidResultLength";
        case FVAD_ERR_INVALID_BINS_LENGTH: return "InvalidBinsLength";
        case FVAD_ERR_OUT_OF_RANGE: return "OutOfRange";
        case FVAD_ERR_NEGATIVE_FREQUENCY: return "NegativeFrequency";
        case FVAD_ERR_CHANNEL_COUNT_MISMATCH: return "ChannelCountMismatch";
        case FVAD_ERR_OUT_OF_MEMORY: return "OutOfMemory";
        case FVAD_ERR_MODEL: return "OnnxError";
        case FVAD_ERR_CUDA: return "CudaError";
        case FVAD_ERR_INVALID_ARGUMENT: return "InvalidArgument";
        case FVAD_ERR_NO_DEVICE: return "NoDevice";
        default: return "Unknown";
    }
}

int fvad_device_count(int *count) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; }
    if (count) *count = n;
    return FVAD_OK;
}

void fvad_vad_machine_config_default(fvad_vad_machine_config *c) {
    if (!c) return;
    c->speech_min_freq = 500; c->speech_max_freq = 2000; c->long_term_speech_avg_sec = 180;
    c->has_initial_long_term_avg = 1; c->initial_long_term_avg = 0.005;
    c->short_term_speech_avg_sec = 0.2f; c->speech_threshold_factor = 10;
    c->channel_vol_ratio_avg_sec = 0.5f; c->channel_vol_ratio_threshold = 0.5f;
    c->min_consecutive_sec_to_open = 0.2f; c->max_speech_gap_sec = 2; c->min_vad_duration_sec = 0.7f;
}

void fvad_config_default(fvad_config *c, size_t sample_rate, size_t n_channels) {
    if (!c) return;
    memset(c, 0, sizeof(*c));
    c->sample_rate = sample_rate; c->n_channels = n_channels;
    c->vad_config.fft_size = 1024;
    fvad_vad_machine_config_default(&c->vad_config.vad_machine_config);
}

void fvad_batch_options_default(fvad_batch_options *o) {
    if (!o) return;
    memset(o, 0, sizeof(*o));
    o->nn_mode = FVAD_NN_F32;
}

/* ------------------------------------------------------------------ FFT operator */

int fvad_fft_create(size_t n_fft, size_t sample_rate, int inverse, fvad_fft **out) {
    return guarded([&] {
        if (!out) throw Error(FVAD_ERR_INVALID_ARGUMENT, "out is NULL");
        *out = nullptr;
        if (n_fft == 0 || (n_fft & 1)) throw Error(FVAD_ERR_INVALID_FFT_SIZE, "n_fft must be even and non-zero");
        require_device();
        auto *h = new fvad_fft;
        h->n_fft = n_fft; h->sample_rate = sample_rate; h->inverse = inverse;
        try { h->plan = new FftPlan(n_fft); } catch (...) { delete h; throw; }
        *out = h;
    });
}

void fvad_fft_destroy(fvad_fft *h) {
    if (!h) return;
    delete h->plan;
    delete h;
}

int fvad_fft_forward_batch(fvad_fft *h, const float *frames, size_t nf, size_t stride, const float *window,
                           fvad_complex *bins, int on_device) {
    return guarded([&] {
        if (!h || !frames || !bins) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
        const size_t n = h->n_fft, nb = n / 2 + 1;
        if (on_device) {
            h->plan->forward(frames, nf, stride, window, bins, 0);
            FVAD_CUDA_CHECK(cudaStreamSynchronize(0));
            return;
        }
        if (!nf) return;
        const size_t span = (nf - 1) * stride + n;
        h->d_real.ensure(span); h->d_bins.ensure(nf * nb);
        FVAD_CUDA_CHECK(cudaMemcpy(h->d_real.p, frames, span * sizeof(float), cudaMemcpyHostToDevice));
        const float *dw = nullptr;
        if (window) {
            h->d_win.ensure(n);
            FVAD_CUDA_CHECK(cudaMemcpy(h->d_win.p, window, n * sizeof(float), cudaMemcpyHostToDevice));
            dw = h->d_win.p;
        }
        h->plan->forward(h->d_real.p, nf, stride, dw, h->d_bins.p, 0);
        FVAD_CUDA_CHECK(cudaMemcpy(bins, h->d_bins.p, nf * nb * sizeof(fvad_complex), cudaMemcpyDeviceToHost));
    });
}

int fvad_fft_inverse_batch(fvad_fft *h, const fvad_complex *bins, size_t nf, float *result, int on_device) {
    return guarded([&] {
        if (!h || !bins || !result) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
        const size_t n = h->n_fft, nb = n / 2 + 1;
        if (on_device) {
            h->plan->inverse(bins, nf, result, 0);
            FVAD_CUDA_CHECK(cudaStreamSynchronize(0));
            return;
        }
        if (!nf) return;
        h->d_real.ensure(nf * n); h->d_bins.ensure(nf * nb);
        FVAD_CUDA_CHECK(cudaMemcpy(h->d_bins.p, bins, nf * nb * sizeof(fvad_complex), cudaMemcpyHostToDevice));
        h->plan->inverse(h->d_bins.p, nf, h->d_real.p, 0);
        FVAD_CUDA_CHECK(cudaMemcpy(result, h->d_real.p, nf * n * sizeof(float), cudaMemcpyDeviceToHost));
    });
}

int fvad_fft_forward(fvad_fft *h, const float *first, size_t n1, const float *second, size_t n2, const float *window,
                     size_t n_window, fvad_complex *bins, size_t n_bins) {
    if (!h) { set_last_error("NULL handle"); return FVAD_ERR_INVALID_ARGUMENT; }
    // argument checks in the order of FFT.zig:91-102
    if (n1 + n2 != h->n_fft) { set_last_error("samples length != n_fft"); return FVAD_ERR_INVALID_SAMPLES_LENGTH; }
    if (n_window != h->n_fft) { set_last_error("window length != n_fft"); return FVAD_ERR_INVALID_WINDOW_LENGTH; }
    if (n_bins != h->n_fft / 2 + 1) { set_last_error("bins length != n_fft/2+1"); return FVAD_ERR_INVALID_RESULT_LENGTH; }
    std::vector<float> joined(h->n_fft);
    if (n1) memcpy(joined.data(), first, n1 * sizeof(float));
    if (n2) memcpy(joined.data() + n1, second, n2 * sizeof(float));
    return fvad_fft_forward_batch(h, joined.data(), 1, h->n_fft, window, bins, 0);
}

int fvad_fft_inverse(fvad_fft *h, const fvad_complex *bins, size_t n_bins, float *result, size_t n_result) {
    if (!h) { set_last_error("NULL handle"); return FVAD_ERR_INVALID_ARGUMENT; }
    if (n_bins != h->n_fft / 2 + 1) { set_last_error("bins length != n_fft/2+1"); return FVAD_ERR_INVALID_BINS_LENGTH; }
    if (n_result != h->n_fft) { set_last_error("result length != n_fft"); return FVAD_ERR_INVALID_RESULT_LENGTH; }
    return fvad_fft_inverse_batch(h, bins, 1, result, 0);
}

size_t fvad_fft_bin_count(const fvad_fft *h) { return h ? h->n_fft / 2 + 1 : 0; }
float fvad_fft_bin_width(const fvad_fft *h) { return h ? (float)h->sample_rate / (float)h->n_fft : 0.0f; }
float fvad_fft_nyquist_freq(const fvad_fft *h) { return h ? (float)h->sample_rate / 2 : 0.0f; }

int fvad_fft_freq_to_bin(const fvad_fft *h, float freq, size_t *bin) {
    if (!h || !bin) { set_last_error("NULL argument"); return FVAD_ERR_INVALID_ARGUMENT; }
    if (freq > fvad_fft_nyquist_freq(h)) { set_last_error("frequency above Nyquist"); return FVAD_ERR_OUT_OF_RANGE; }
    if (freq < 0) { set_last_error("negative frequency"); return FVAD_ERR_NEGATIVE_FREQUENCY; }
    *bin = (size_t)roundf(freq / fvad_fft_bin_width(h));
    return FVAD_OK;
}

int fvad_fft_bin_to_freq(const fvad_fft *h, size_t bin, float *freq) {
    if (!h || !freq) { set_last_error("NULL argument"); return FVAD_ERR_INVALID_ARGUMENT; }
    if (bin > fvad_fft_bin_count(h) - 1) { set_last_error("bin index out of range"); return FVAD_ERR_OUT_OF_RANGE; }
    *freq = (float)bin * fvad_fft_bin_width(h);
    return FVAD_OK;
}

}  // extern "C"

/* ------------------------------------------------------------------ helpers for the handles below */
namespace {

const char *kDefaultModel = "data/nsnet2-20ms-baseline.onnx";  // src/NSNet2.zig:56

Batch *make_batch(size_t n_streams, const fvad_config &cfg, const fvad_nsnet2_weights *w, const fvad_batch_options &opt) {
    require_device();
    HostWeights hw;
    const HostWeights *hwp = nullptr;
    const bool needs_nn = !cfg.skip_processing && !opt.stft_only && !opt.forced_gains;
    if (needs_nn) {
        if (w) hw.from_struct(*w);
        else hw.from_file(cfg.vad_config.denoiser_model_path ? cfg.vad_config.denoiser_model_path : kDefaultModel);
        hwp = &hw;
    }
    return new Batch(n_streams, cfg, hwp, opt);
}

}  // namespace

extern "C" {

/* ------------------------------------------------------------------ NSNet2 operator */

size_t fvad_nsnet2_chunk_size(size_t sample_rate) {
    // NSNet2.getChunkSize, src/NSNet2.zig:157-159 (calcDownsampleRate panics when not divisible; 0 here)
    if (sample_rate % 16000 != 0) return 0;
    return (size_t)kChunk16 * (sample_rate / 16000);
}

int fvad_nsnet2_create_with_weights(size_t sample_rate, const fvad_nsnet2_weights *w, int nn_mode, fvad_nsnet2 **out) {
    return guarded([&] {
        if (!out) throw Error(FVAD_ERR_INVALID_ARGUMENT, "out is NULL");
        *out = nullptr;
        fvad_config cfg;
        fvad_config_default(&cfg, sample_rate, 1);
        fvad_batch_options opt;
        fvad_batch_options_default(&opt);
        opt.nn_mode = nn_mode;
        opt.keep_taps = 1;
        auto *h = new fvad_nsnet2;
        try { h->impl = make_batch(1, cfg, w, opt); } catch (...) { delete h; throw; }
        h->sample_rate = sample_rate;
        *out = h;
    });
}

int fvad_nsnet2_create(size_t sample_rate, const char *model_path, fvad_nsnet2 **out) {
    return guarded([&] {
        if (!out) throw Error(FVAD_ERR_INVALID_ARGUMENT, "out is NULL");
        *out = nullptr;
        HostWeights hw;
        hw.from_file(model_path ? model_path : kDefaultModel);
        fvad_nsnet2_weights w;
        memset(&w, 0, sizeof(w));
        w.n_bins = hw.n_bins; w.hidden = hw.hidden; w.fc2 = hw.fc2; w.fc3 = hw.fc3; w.fc1_relu = hw.fc1_relu;
        w.fc1_w = hw.fc1_w.data(); w.fc1_b = hw.fc1_b.data();
        w.gru1_W = hw.gru_W[0].data(); w.gru1_R = hw.gru_R[0].data(); w.gru1_Wb = hw.gru_Wb[0].data(); w.gru1_Rb = hw.gru_Rb[0].data();
        w.gru2_W = hw.gru_W[1].data(); w.gru2_R = hw.gru_R[1].data(); w.gru2_Wb = hw.gru_Wb[1].data(); w.gru2_Rb = hw.gru_Rb[1].data();
        w.fc2_w = hw.fc2_w.data(); w.fc2_b = hw.fc2_b.data(); w.fc3_w = hw.fc3_w.data(); w.fc3_b = hw.fc3_b.data();
        w.fc4_w = hw.fc4_w.data(); w.fc4_b = hw.fc4_b.data();
        const int rc = fvad_nsnet2_create_with_weights(sample_rate, &w, FVAD_NN_F32, out);
        if (rc != FVAD_OK) throw Error(rc, last_error());
    });
}

void fvad_nsnet2_destroy(fvad_nsnet2 *h) {
    if (!h) return;
    delete h->impl;
    delete h;
}

int fvad_nsnet2_denoise(fvad_nsnet2 *h, const float *first, size_t n1, const float *second, size_t n2, float *out, size_t n_out) {
    return guarded([&] {
        if (!h || !out) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
        const size_t chunk = h->impl->chunk();
        if (n1 + n2 != chunk) throw Error(FVAD_ERR_INVALID_INPUT_LENGTH, "input length != chunk size");  // NSNet2.zig:167-169
        if (n_out != chunk) throw Error(FVAD_ERR_INVALID_RESULT_LENGTH, "result length != chunk size");
        h->joined.resize(chunk);
        if (n1) memcpy(h->joined.data(), first, n1 * sizeof(float));
        if (n2) memcpy(h->joined.data() + n1, second, n2 * sizeof(float));
        const float *p = h->joined.data();
        h->impl->push(&p, 1, chunk, 0);
        h->impl->tap(FVAD_TAP_DENOISED, out, n_out * sizeof(float));
    });
}

int fvad_nsnet2_calc_spectrogram(const float *audio, size_t n_audio, size_t n_frames, const float *window, size_t n_window,
                                 fvad_complex *result, size_t n_result) {
    return guarded([&] {
        if (!audio || !window || !result) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
        if (n_window != (size_t)kNfft) throw Error(FVAD_ERR_INVALID_WINDOW_LENGTH, "window length != 320");
        if (n_audio < n_frames * kHop + kHop) throw Error(FVAD_ERR_INVALID_SAMPLES_LENGTH, "audio chunk too short for n_frames");
        if (n_result != n_frames * kBins) throw Error(FVAD_ERR_INVALID_RESULT_LENGTH, "result length != n_frames*161");
        fvad_fft *f = nullptr;
        int rc = fvad_fft_create(kNfft, 16000, 0, &f);
        if (rc != FVAD_OK) throw Error(rc, last_error());
        rc = fvad_fft_forward_batch(f, audio, n_frames, kHop, window, result, 0);
        fvad_fft_destroy(f);
        if (rc != FVAD_OK) throw Error(rc, last_error());
    });
}

/* ------------------------------------------------------------------ pipeline facade */

int fvad_pipeline_create_with_weights(const fvad_config *cfg, const fvad_callbacks *cb, const fvad_nsnet2_weights *w, int nn_mode,
                                      fvad_pipeline **out) {
    return guarded([&] {
        if (!cfg || !out) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
        *out = nullptr;
        fvad_batch_options opt;
        fvad_batch_options_default(&opt);
        opt.nn_mode = nn_mode;
        auto *p = new fvad_pipeline;
        if (cb && (cb->on_original_recording || cb->on_denoised_recording)) { p->cb = *cb; p->has_cb = true; opt.keep_taps = 1; }
        try { p->impl = make_batch(1, *cfg, w, opt); } catch (...) { delete p; throw; }
        p->sample_rate = cfg->sample_rate;
        p->n_channels = cfg->n_channels;
        p->orig.ch.resize(cfg->n_channels);
        p->den.ch.resize(cfg->n_channels);
        *out = p;
    });
}

int fvad_pipeline_create(const fvad_config *cfg, const fvad_callbacks *cb, fvad_pipeline **out) {
    return fvad_pipeline_create_with_weights(cfg, cb, nullptr, FVAD_NN_F32, out);
}

void fvad_pipeline_destroy(fvad_pipeline *p) {
    if (!p) return;
    delete p->impl;
    delete p;
}

int fvad_pipeline_push_samples(fvad_pipeline *p, const float *const *pcm, size_t n_channels, size_t n, uint64_t *first) {
    return guarded([&] {
        if (!p) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL handle");
        if (first) *first = p->impl->total_write_count();  // AudioPipeline.zig:119
        p->impl->push(pcm, n_channels, n, 0);
        if (p->has_cb && !p->impl->skipping()) run_recorders(p, pcm, n);
    });
}

uint64_t fvad_pipeline_total_write_count(const fvad_pipeline *p) { return p ? p->impl->total_write_count() : 0; }

int fvad_pipeline_segments(fvad_pipeline *p, size_t machine, const fvad_segment **segs, size_t *n) {
    return guarded([&] {
        if (!p) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL handle");
        p->impl->segments(0, machine, segs, n);
    });
}

/* ------------------------------------------------------------------ batch engine */

int fvad_batch_create(size_t n_streams, const fvad_config *cfg, const fvad_nsnet2_weights *w, const fvad_batch_options *opt,
                      fvad_batch **out) {
    return guarded([&] {
        if (!cfg || !out) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
        *out = nullptr;
        fvad_batch_options o;
        if (opt) o = *opt; else fvad_batch_options_default(&o);
        auto *b = new fvad_batch;
        try { b->impl = make_batch(n_streams, *cfg, w, o); } catch (...) { delete b; throw; }
        *out = b;
    });
}

void fvad_batch_destroy(fvad_batch *b) {
    if (!b) return;
    delete b->impl;
    delete b;
}

int fvad_batch_push(fvad_batch *b, const float *const *pcm, size_t n_pointers, size_t n, int memory) {
    return guarded([&] {
        if (!b) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL handle");
        b->impl->push(pcm, n_pointers, n, memory);
    });
}

int fvad_batch_push_i16(fvad_batch *b, const int16_t *const *pcm, size_t n_pointers, size_t n, int memory) {
    return guarded([&] {
        if (!b) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL handle");
        b->impl->push_i16(pcm, n_pointers, n, memory);
    });
}

int fvad_batch_segments(fvad_batch *b, size_t stream, size_t machine, const fvad_segment **segs, size_t *n) {
    return guarded([&] {
        if (!b) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL handle");
        b->impl->segments(stream, machine, segs, n);
    });
}

int fvad_batch_tap(fvad_batch *b, int tap, void *dst, size_t dst_bytes, size_t *n_elements) {
    return guarded([&] {
        if (!b) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL handle");
        const size_t n = b->impl->tap(tap, dst, dst_bytes);
        if (n_elements) *n_elements = n;
    });
}

int fvad_batch_stats_get(fvad_batch *b, fvad_batch_stats *out) {
    return guarded([&] {
        if (!b || !out) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
        b->impl->stats(out);
    });
}

void fvad_eval_config_default(fvad_eval_config *cfg) {
    if (!cfg) return;
    cfg->ignore_shorter_than_sec = 0.0f;
    cfg->extrude_start = 0.0f;
    cfg->extrude_end = 0.0f;
    cfg->fill_gaps = 0.0f;
    cfg->f_score_beta = 0.7f;  // statistics.zig:108
}

int fvad_evaluate(const fvad_eval_segment *vad, size_t n_vad, const fvad_eval_segment *ref, size_t n_ref,
                  const fvad_eval_config *cfg, fvad_eval_stats *out) {
    return guarded([&] {
        fvad_eval_config c;
        fvad_eval_config_default(&c);
        evaluate_lists(vad, n_vad, ref, n_ref, cfg ? *cfg : c, out);
    });
}

int fvad_vad_machine_scan(const fvad_vad_machine_config *cfgs, size_t n_machines, size_t sample_rate, size_t fft_size,
                          const float *band_min, const float *ratio, size_t n_frames, size_t frames_per_launch, uint8_t *met,
                          fvad_segment *segments, size_t max_segments, size_t *n_segments) {
    return guarded([&] {
        if (fft_size == 0 || sample_rate == 0) throw Error(FVAD_ERR_INVALID_ARGUMENT, "sample_rate and fft_size must be positive");
        vad_machine_scan(cfgs, n_machines, sample_rate, fft_size, band_min, ratio, n_frames, frames_per_launch, met, segments,
                         max_segments, n_segments);
    });
}

int fvad_batch_evaluate(fvad_batch *b, const fvad_eval_segment *refs, const size_t *ref_offsets, const fvad_eval_config *cfg,
                        fvad_eval_stats *out) {
    return guarded([&] {
        if (!b) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL handle");
        fvad_eval_config c;
        fvad_eval_config_default(&c);
        b->impl->evaluate(refs, ref_offsets, cfg ? *cfg : c, out);
    });
}

}  // extern "C"
