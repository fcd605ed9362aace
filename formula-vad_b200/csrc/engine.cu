// engine.cu -- see engine.cuh.
#include "engine.cuh"
#include "ops.cuh"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace fvad {

namespace {

// FFT.freqToBin, reference src/FFT.zig:140-167 (f32 arithmetic)
int freq_to_bin(float freq, size_t sample_rate, size_t n_fft) {
    const float nyq = (float)sample_rate / 2;
    if (freq > nyq) throw Error(FVAD_ERR_OUT_OF_RANGE, "speech band above Nyquist");
    if (freq < 0) throw Error(FVAD_ERR_NEGATIVE_FREQUENCY, "negative speech band frequency");
    return (int)roundf(freq / ((float)sample_rate / (float)n_fft));
}

__global__ void init_scan_kernel(ScanState *states, double *rings, const ScanMachine *machines, int S, int M, int pitch) {
    const int t = blockIdx.x;  // (stream, machine)
    if (t >= S * M) return;
    const ScanMachine mc = machines[t % M];
    double *ring = rings + (size_t)t * pitch;
    for (int i = threadIdx.x; i < pitch; i += blockDim.x) ring[i] = (mc.has_init && i < mc.n_long) ? mc.init_avg : 0.0;
    if (threadIdx.x == 0) {
        ScanState st;
        memset(&st, 0, sizeof(st));
        if (mc.has_init) {
            // RollingAverage.init pre-fill (reference src/structures/RollingAverage.zig:19-25):
            // written_count = len and last_avg = sum_i init*(1/len), accumulated sequentially in f64
            st.long_cnt = mc.n_long;
            const double scalar = 1.0 / (double)mc.n_long;
            double acc = 0.0, sum = 0.0;
            for (int i = 0; i < mc.n_long; ++i) { acc += mc.init_avg * scalar; sum += mc.init_avg; }
            st.last_avg = acc;
            st.long_sum = sum;
            st.has_last = 1;
        }
        states[t] = st;
    }
}

}  // namespace

// VADMachine.init, reference src/AudioPipeline/VADMachine.zig:75-106 (f32 products, truncation)
ScanMachine derive_machine(const fvad_vad_machine_config &c, size_t sample_rate, size_t fft_size) {
    ScanMachine d;
    memset(&d, 0, sizeof(d));
    const float sr_f = (float)sample_rate, fft_f = (float)fft_size;
    const float evals = sr_f / fft_f;
    d.n_long = std::max<long long>(1, (long long)(evals * c.long_term_speech_avg_sec));
    d.n_short = std::max<long long>(1, (long long)(evals * c.short_term_speech_avg_sec));
    d.n_ratio = std::max<long long>(1, (long long)(evals * c.channel_vol_ratio_avg_sec));  // reference: no max(1); 0 would trap
    d.has_init = c.has_initial_long_term_avg;
    d.init_avg = c.initial_long_term_avg;
    d.threshold_factor = (double)c.speech_threshold_factor;
    d.ratio_threshold = (double)c.channel_vol_ratio_threshold;
    // VADMachine.run :160-163, getOffsetRecording{Start,End} :312-325
    d.min_to_open = (unsigned long long)(sr_f * c.min_consecutive_sec_to_open);
    d.max_gap = (unsigned long long)(sr_f * c.max_speech_gap_sec);
    d.start_margin = d.end_margin = (unsigned long long)(sr_f * 2);
    d.min_duration_sec = c.min_vad_duration_sec;
    d.frame_sec = fft_f / sr_f;  // trackSpeechStats :248-249
    d.sample_rate_f = sr_f;
    return d;
}

// The VAD machine as an operator on a prepared series: every machine scans its own row of band minima.
void vad_machine_scan(const fvad_vad_machine_config *cfgs, size_t n_machines, size_t sample_rate, size_t fft_size,
                      const float *band_min, const float *ratio, size_t n_frames, size_t frames_per_launch,
                      unsigned char *met, fvad_segment *segments, size_t max_segments, size_t *n_segments) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) throw Error(FVAD_ERR_NO_DEVICE, "no CUDA device: libfvad_cuda has no CPU path");
    if (!cfgs || !n_machines || (n_frames && (!band_min || !ratio)) || !n_segments || (max_segments && !segments))
        throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL argument");
    const int M = (int)n_machines;
    std::vector<ScanMachine> hm(n_machines);
    int pitch = 0, fast = 1;
    for (int m = 0; m < M; ++m) {
        hm[(size_t)m] = derive_machine(cfgs[m], sample_rate, fft_size);
        hm[(size_t)m].band = m;
        pitch = std::max(pitch, hm[(size_t)m].n_long + hm[(size_t)m].n_short + hm[(size_t)m].n_ratio);
        if (hm[(size_t)m].n_long < 2 * 256 + 2 || hm[(size_t)m].n_short + hm[(size_t)m].n_ratio > 64) fast = 0;
    }
    const size_t nf = std::max<size_t>(n_frames, 1), ms = std::max<size_t>(max_segments, 1);
    DevBuf<ScanMachine> d_m; DevBuf<ScanState> d_st; DevBuf<double> d_rings; DevBuf<fvad_segment> d_seg;
    DevBuf<float> d_band, d_ratio; DevBuf<unsigned char> d_met;
    d_m.ensure(n_machines); d_st.ensure(n_machines); d_rings.ensure(n_machines * (size_t)pitch); d_seg.ensure(n_machines * ms);
    d_band.ensure(n_machines * nf); d_ratio.ensure(nf); d_met.ensure(n_machines * nf);
    FVAD_CUDA_CHECK(cudaMemcpy(d_m.p, hm.data(), hm.size() * sizeof(ScanMachine), cudaMemcpyHostToDevice));
    if (n_frames) {
        FVAD_CUDA_CHECK(cudaMemcpy(d_band.p, band_min, n_machines * n_frames * sizeof(float), cudaMemcpyHostToDevice));
        FVAD_CUDA_CHECK(cudaMemcpy(d_ratio.p, ratio, n_frames * sizeof(float), cudaMemcpyHostToDevice));
    }
    init_scan_kernel<<<(unsigned)M, 128>>>(d_st.p, d_rings.p, d_m.p, 1, M, pitch);
    FVAD_LAUNCHED();
    const size_t step = frames_per_launch ? frames_per_launch : nf;
    for (size_t f0 = 0; f0 < n_frames; f0 += step) {
        ScanArgs ca;
        ca.band = d_band.p + f0; ca.ratio = d_ratio.p + f0; ca.machines = d_m.p; ca.states = d_st.p;
        ca.rings = d_rings.p; ca.ring_pitch = pitch; ca.segments = d_seg.p; ca.max_segments = (int)max_segments;
        ca.met = d_met.p + f0; ca.met_per_machine = 1; ca.S = 1; ca.M = M;
        ca.n_frames = (int)std::min(step, n_frames - f0); ca.max_frames = (int)nf;
        ca.first_frame = (long long)f0; ca.fft_size = (int)fft_size; ca.fast = fast;
        launch_vadscan(ca, 0);
    }
    FVAD_CUDA_CHECK(cudaGetLastError());
    std::vector<ScanState> hs(n_machines);
    FVAD_CUDA_CHECK(cudaMemcpy(hs.data(), d_st.p, n_machines * sizeof(ScanState), cudaMemcpyDeviceToHost));
    if (met && n_frames) FVAD_CUDA_CHECK(cudaMemcpy(met, d_met.p, n_machines * n_frames, cudaMemcpyDeviceToHost));
    for (int m = 0; m < M; ++m) {
        n_segments[m] = hs[(size_t)m].n_segments;
        const size_t have = std::min<size_t>(hs[(size_t)m].n_segments, max_segments);
        if (have) FVAD_CUDA_CHECK(cudaMemcpy(segments + (size_t)m * max_segments, d_seg.p + (size_t)m * max_segments,
                                             have * sizeof(fvad_segment), cudaMemcpyDeviceToHost));
    }
}

void Batch::init_machines(const fvad_config &cfg) {
    const fvad_vad_config &vc = cfg.vad_config;
    M_ = 1 + (int)(vc.alt_vad_machine_configs ? vc.n_alt_vad_machine_configs : 0);
    machines_.resize((size_t)M_);
    std::vector<int> lo, hi;
    const float sr_f = (float)sample_rate_, fft_f = (float)fft_size_;
    for (int m = 0; m < M_; ++m) {
        const fvad_vad_machine_config &c = m ? vc.alt_vad_machine_configs[m - 1] : vc.vad_machine_config;
        MachineDerived md;
        memset(&md, 0, sizeof(md));
        md.bin_lo = freq_to_bin(c.speech_min_freq, sample_rate_, (size_t)fft_size_);
        md.bin_hi = freq_to_bin(c.speech_max_freq, sample_rate_, (size_t)fft_size_);
        int b = -1;
        for (size_t i = 0; i < lo.size(); ++i)
            if (lo[i] == md.bin_lo && hi[i] == md.bin_hi) b = (int)i;
        if (b < 0) { b = (int)lo.size(); lo.push_back(md.bin_lo); hi.push_back(md.bin_hi); }
        md.dev = derive_machine(c, sample_rate_, (size_t)fft_size_);
        md.dev.band = b;
        const ScanMachine &d = md.dev;
        machines_[(size_t)m] = md;
        ring_pitch_ = std::max(ring_pitch_, d.n_long + d.n_short + d.n_ratio);
        if (d.n_long < 2 * 256 + 2 || d.n_short + d.n_ratio > 64) scan_fast_ = 0;  // vadscan.cu fast-path limits
    }
    n_bands_ = (int)lo.size();
    band_lo_.ensure(lo.size());
    band_hi_.ensure(hi.size());
    FVAD_CUDA_CHECK(cudaMemcpy(band_lo_.p, lo.data(), lo.size() * sizeof(int), cudaMemcpyHostToDevice));
    FVAD_CUDA_CHECK(cudaMemcpy(band_hi_.p, hi.data(), hi.size() * sizeof(int), cudaMemcpyHostToDevice));
    std::vector<ScanMachine> dm((size_t)M_);
    for (int m = 0; m < M_; ++m) dm[(size_t)m] = machines_[(size_t)m].dev;
    d_machines_.ensure((size_t)M_);
    FVAD_CUDA_CHECK(cudaMemcpy(d_machines_.p, dm.data(), dm.size() * sizeof(ScanMachine), cudaMemcpyHostToDevice));
    const size_t SM = (size_t)S_ * M_;
    d_states_.ensure(SM);
    rings_.ensure(SM * (size_t)ring_pitch_);
    max_segments_ = (SM > 64) ? 1024 : 4096;
    d_segments_.ensure(SM * (size_t)max_segments_);
    init_scan_kernel<<<(unsigned)SM, 128, 0, stream_>>>(d_states_.p, rings_.p, d_machines_.p, S_, M_, ring_pitch_);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
    h_states_.assign(SM, ScanState{});
    h_segments_.resize(SM * (size_t)max_segments_);
}

Batch::Batch(size_t n_streams, const fvad_config &cfg, const HostWeights *weights, const fvad_batch_options &opt)
    : S_((int)n_streams), C_((int)cfg.n_channels), SC_((int)(n_streams * cfg.n_channels)),
      sample_rate_(cfg.sample_rate), opt_(opt) {
    // VADPipeline.init rejects anything but 48 kHz (reference VADPipeline.zig:55-58).  Documented
    // deviation: 16 kHz is also accepted, as the same code with downsample rate 1.
    if (sample_rate_ != 48000 && sample_rate_ != 16000)
        throw Error(FVAD_ERR_INVALID_SAMPLE_RATE, "sample_rate must be 48000 (or 16000, documented deviation)");
    if (S_ <= 0 || C_ <= 0) throw Error(FVAD_ERR_INVALID_ARGUMENT, "n_streams and n_channels must be positive");
    fft_size_ = (int)cfg.vad_config.fft_size;
    if (fft_size_ != 1024) throw Error(FVAD_ERR_INVALID_FFT_SIZE, "vad_config.fft_size: the GPU path implements 1024");
    D_ = (int)(sample_rate_ / 16000);
    skip_ = cfg.skip_processing != 0;
    chunk_ = kChunk16 * D_;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        throw Error(FVAD_ERR_NO_DEVICE, "no CUDA device: libfvad_cuda has no CPU path");
    if (opt_.device < 0 || opt_.device >= ndev) throw Error(FVAD_ERR_INVALID_ARGUMENT, "bad device ordinal");
    FVAD_CUDA_CHECK(cudaSetDevice(opt_.device));
    FVAD_CUDA_CHECK(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
    if (getenv("FVAD_SCAN_SERIAL")) scan_stream_ = stream_;  // experiments: scan on the main stream (no overlap)
    else FVAD_CUDA_CHECK(cudaStreamCreateWithFlags(&scan_stream_, cudaStreamNonBlocking));
    FVAD_CUDA_CHECK(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
        FVAD_CUDA_CHECK(cudaEventCreateWithFlags(&stage_ready_[i], cudaEventDisableTiming));
        FVAD_CUDA_CHECK(cudaEventCreateWithFlags(&stage_free_[i], cudaEventDisableTiming));
    }
    for (auto &e : ev_) FVAD_CUDA_CHECK(cudaEventCreate(&e));
    FVAD_CUDA_CHECK(cudaEventCreateWithFlags(&scan_done_, cudaEventDisableTiming));
    FVAD_CUDA_CHECK(cudaEventRecord(scan_done_, scan_stream_));
    for (auto &e : scan_done_b_) {
        FVAD_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        FVAD_CUDA_CHECK(cudaEventRecord(e, scan_stream_));
    }
    FVAD_CUDA_CHECK(cudaEventCreateWithFlags(&after_rnn_, cudaEventDisableTiming));
    defer_scan_ = scan_stream_ != stream_ && !opt_.keep_taps && !getenv("FVAD_SCAN_EAGER");
    tables_.build();
    launches0_ = g_kernel_launches;

    // span size: bounded by memory for the per-frame intermediates (~3.2 KB per frame-row)
    // auto: two full waves of 128-sequence GRU tiles (one CTA per SM owns a tile for all 54 steps), so the
    // recurrent kernel never runs a partial wave except in a push's last span
    int n_sm = 148;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, opt_.device);
    size_t kmax = opt_.max_span_chunks ? opt_.max_span_chunks : std::max<size_t>(1, (size_t)n_sm * 2 * 128 / (size_t)SC_);
    Kmax_ = (int)std::min<size_t>(kmax, 16384);

    if (!cfg.skip_processing && !opt_.stft_only && !opt_.forced_gains) {
        if (!weights) throw Error(FVAD_ERR_MODEL, "no NSNet2 weights given");
        switch (opt_.nn_mode) {
            case FVAD_NN_F32: nn_ = make_nn_f32(*weights); break;
            case FVAD_NN_TC_F16X3: nn_ = make_nn_tc(*weights, true); break;
            case FVAD_NN_TC_F16: nn_ = make_nn_tc(*weights, false); break;
            default: throw Error(FVAD_ERR_INVALID_ARGUMENT, "unknown nn_mode");
        }
    }
    if (opt_.forced_gains) {
        const size_t n = (size_t)SC_ * opt_.forced_gain_frames * kBins;
        forced_.ensure(n);
        FVAD_CUDA_CHECK(cudaMemcpy(forced_.p, opt_.forced_gains, n * sizeof(float), cudaMemcpyHostToDevice));
    }
    for (int i = 0; i < 2; ++i) {
        halo_[i].ensure((size_t)SC_ * kHop); halo_[i].zero(stream_);
        tail_[i].ensure((size_t)SC_ * kHop); tail_[i].zero(stream_);
    }
    ctx_carry_.ensure((size_t)SC_ * kCtx * kBins); ctx_carry_.zero(stream_);  // NSNet2.zig:79 context rows start at 0.0
    ctx16_carry_.ensure((size_t)SC_ * kCtx * 88); ctx16_carry_.zero(stream_);
    ylook_carry_.ensure((size_t)SC_ * kYLook); ylook_carry_.zero(stream_);
    sumsq_carry_.ensure((size_t)SC_); sumsq_carry_.zero(stream_);
    pend_.ensure((size_t)SC_ * chunk_);
    d_xptr_.ensure((size_t)SC_);
    init_machines(cfg);
    FVAD_CUDA_CHECK(cudaStreamSynchronize(stream_));
}

Batch::~Batch() {
    if (stream_) cudaStreamSynchronize(stream_);
    if (scan_stream_ && scan_stream_ != stream_) { cudaStreamSynchronize(scan_stream_); cudaStreamDestroy(scan_stream_); }
    if (scan_done_) cudaEventDestroy(scan_done_);
    for (auto &e : scan_done_b_) if (e) cudaEventDestroy(e);
    if (after_rnn_) cudaEventDestroy(after_rnn_);
    if (copy_stream_) { cudaStreamSynchronize(copy_stream_); cudaStreamDestroy(copy_stream_); }
    for (int i = 0; i < 2; ++i) {
        if (stage_ready_[i]) cudaEventDestroy(stage_ready_[i]);
        if (stage_free_[i]) cudaEventDestroy(stage_free_[i]);
    }
    tables_.release();
    for (auto &e : ev_) if (e) cudaEventDestroy(e);
    for (auto &sp : span_ev_) for (auto &e : sp) cudaEventDestroy(e);
    if (stream_) cudaStreamDestroy(stream_);
}


// One pass over K whole chunks of every stream-channel.  d_x: device array of SC pointers; the
// span starts x_offset samples into each.
void Batch::process_span(const float *const *d_x, size_t x_offset, int K) {
    const size_t rows = (size_t)kFrames * K;        // frame rows per stream-channel
    const size_t frow = (size_t)kCtx + rows;
    const size_t ypitch = (size_t)kYLook + (size_t)kChunk16 * K;
    const bool full = !opt_.stft_only;
    spec_.ensure((size_t)SC_ * rows * kBins);
    // feature rows: fp16 for a backend that reads them so, f32 for the others and for the taps
    const bool want16 = nn_ && full && !forced_.p && nn_->wants_feat16();
    const bool want32 = !want16 || opt_.keep_taps;
    if (want32) feat_.ensure((size_t)SC_ * frow * kBins);
    if (want16) feat16_.ensure((size_t)SC_ * frow * 88);
    sumsq_.ensure((size_t)SC_ * (K + 1));
    if (full) {
        gains_.ensure((size_t)SC_ * rows * kBins);
        y16_.ensure((size_t)SC_ * ypitch);
    }
    cudaStream_t s = stream_;

    // per-span stage events (summed over the push's spans after the final synchronise)
    if (span_idx_ >= span_ev_.size()) {
        std::array<cudaEvent_t, 7> e{};
        for (auto &x : e) FVAD_CUDA_CHECK(cudaEventCreate(&x));
        span_ev_.push_back(e);
    }
    cudaEvent_t *ev_ = span_ev_[span_idx_++].data();

    // ---- K1 ----
    FVAD_CUDA_CHECK(cudaEventRecord(ev_[0], s));
    StftArgs sa;
    sa.x = d_x; sa.x_offset = x_offset;
    sa.halo_in = halo_[parity_].p; sa.halo_out = halo_[parity_ ^ 1].p;
    sa.spec = spec_.p; sa.feat = want32 ? feat_.p : nullptr; sa.sumsq = sumsq_.p;
    sa.K = K; sa.SC = SC_; sa.D = D_;
    sa.feat16 = want16 ? feat16_.p : nullptr;
    launch_stft(sa, tables_, s);
    // 4 context feature rows carried from the previous span (NSNet2.zig:188-203)
    if (want32) launch_copy_rows(feat_.p, ctx_carry_.p, frow * kBins, (size_t)kCtx * kBins, (size_t)kCtx * kBins, (size_t)SC_, s);
    if (want16) launch_copy_rows(feat16_.p, ctx16_carry_.p, frow * 88, (size_t)kCtx * 88, (size_t)kCtx * 88, (size_t)SC_, s);
    FVAD_CUDA_CHECK(cudaEventRecord(ev_[1], s));

    const uint64_t span_start = processed_;
    const uint64_t span_end = processed_ + (uint64_t)K * chunk_;
    int n_frames = 0;
    if (full) {
        // ---- NSNet2 ----
        if (forced_.p) {
            const size_t f0 = (size_t)(processed_ / (uint64_t)chunk_) * kFrames;
            if (f0 + rows > opt_.forced_gain_frames) throw Error(FVAD_ERR_INVALID_ARGUMENT, "forced_gains shorter than the audio");
            launch_copy_rows(gains_.p, forced_.p + f0 * kBins, rows * kBins, opt_.forced_gain_frames * kBins, rows * kBins, (size_t)SC_, s);
            if (pscan_.valid) issue_scan(nullptr);
        } else {
            NnSpan ns{want32 ? feat_.p : nullptr, gains_.p, SC_, K, want16 ? feat16_.p : nullptr};
            // The previous span's VAD scan starts once this span's recurrent layers are queued: their epilogue
            // warps are issue-bound, and a scan warp sharing an SM sub-partition with them stretched the whole
            // GRU launch by a third, while the GEMMs and the spectral kernels that follow do not notice it.
            if (pscan_.valid) ns.after_rnn = after_rnn_;
            nn_->forward(ns, s);
            if (pscan_.valid) issue_scan(after_rnn_);
        }
        FVAD_CUDA_CHECK(cudaEventRecord(ev_[2], s));

        // ---- K3a ----
        IstftArgs ia;
        ia.spec = spec_.p; ia.gains = gains_.p;
        ia.tail_in = tail_[parity_].p; ia.tail_out = tail_[parity_ ^ 1].p;
        ia.y16 = y16_.p; ia.K = K; ia.SC = SC_;
        launch_istft(ia, tables_, s);
        launch_copy_rows(y16_.p, ylook_carry_.p, ypitch, kYLook, kYLook, (size_t)SC_, s);
        launch_copy_rows(sumsq_.p, sumsq_carry_.p, (size_t)K + 1, 1, 1, (size_t)SC_, s);
        FVAD_CUDA_CHECK(cudaEventRecord(ev_[3], s));

        // ---- K3b ----  (band_/ratio_/met_[vb] were last read by the scan two spans back)
        const int vb = vbuf_;
        FVAD_CUDA_CHECK(cudaStreamWaitEvent(s, scan_done_b_[vb], 0));
        const uint64_t frames_end = span_end / (uint64_t)fft_size_;
        n_frames = (int)(frames_end - frames_done_);
        if (n_frames > 0) {
            if (n_frames > max_frames_) {
                if (pscan_.valid) issue_scan(nullptr);
                FVAD_CUDA_CHECK(cudaStreamSynchronize(scan_stream_));  // earlier scans still read the old buffers
                max_frames_ = n_frames + 64;
                for (int b = 0; b < 2; ++b) {
                    band_[b].ensure((size_t)n_bands_ * S_ * max_frames_);
                    ratio_[b].ensure((size_t)S_ * max_frames_);
                    met_[b].ensure((size_t)S_ * max_frames_);
                }
            }
            VadFftArgs va;
            va.y16 = y16_.p; va.sumsq = sumsq_.p; va.band = band_[vb].p; va.ratio = ratio_[vb].p;
            va.band_lo = band_lo_.p; va.band_hi = band_hi_.p; va.n_bands = n_bands_;
            va.S = S_; va.C = C_; va.K = K; va.D = D_;
            va.span_start = (long long)span_start; va.first_frame = (long long)frames_done_;
            va.n_frames = n_frames; va.max_frames = max_frames_; va.fft_size = fft_size_; va.chunk = chunk_;
            va.has_prev_chunk = span_start > 0;
            work_ctr_.ensure(4);
            va.work_ctr = work_ctr_.p;
            launch_vadfft(va, tables_, s);
        }
        FVAD_CUDA_CHECK(cudaEventRecord(ev_[4], s));

        // ---- K4 ----  on its own stream: a latency-bound scan that overlaps the next span's kernels
        if (pscan_.valid) issue_scan(nullptr);  // (only when the span above had no network launch)
        pscan_.valid = true; pscan_.buf = vb; pscan_.n_frames = n_frames;
        pscan_.first_frame = (long long)frames_done_; pscan_.span = span_idx_ - 1;
        if (n_frames > 0) vbuf_ ^= 1;
        if (!defer_scan_ || n_frames <= 0) issue_scan(nullptr);
        if (opt_.keep_taps) FVAD_CUDA_CHECK(cudaStreamWaitEvent(s, scan_done_, 0));  // the taps below copy met_
    }

    // ---- taps ----
    if (opt_.keep_taps && taps_.fill_chunks + K <= taps_.cap_chunks) {
        const size_t trows = (size_t)kFrames * taps_.cap_chunks, off = (size_t)kFrames * taps_.fill_chunks;
        launch_copy_rows(reinterpret_cast<float *>(taps_.spec.p) + off * kBins * 2, reinterpret_cast<const float *>(spec_.p),
                         trows * kBins * 2, rows * kBins * 2, rows * kBins * 2, (size_t)SC_, s);
        launch_copy_rows(taps_.feat.p + off * kBins, feat_.p + (size_t)kCtx * kBins, trows * kBins, frow * kBins, rows * kBins, (size_t)SC_, s);
        if (full) {
            launch_copy_rows(taps_.gains.p + off * kBins, gains_.p, trows * kBins, rows * kBins, rows * kBins, (size_t)SC_, s);
            launch_copy_rows(taps_.y16.p + (size_t)kChunk16 * taps_.fill_chunks, y16_.p + kYLook, (size_t)kChunk16 * taps_.cap_chunks,
                             ypitch, (size_t)kChunk16 * K, (size_t)SC_, s);
            // denoised PCM at the input rate (what AudioPipeline.pushDenoisedSamples receives, AudioPipeline.zig:145-167)
            launch_upsample(y16_.p + kYLook, ypitch, taps_.den.p + (size_t)chunk_ * taps_.fill_chunks, (size_t)chunk_ * taps_.cap_chunks,
                            (size_t)kChunk16 * K, D_, SC_, s);
            if (n_frames > 0 && taps_.fill_vad + n_frames <= taps_.cap_vad) {
                launch_copy_rows(taps_.band.p + taps_.fill_vad, band_[vbuf_ ^ 1].p, taps_.cap_vad, max_frames_, (size_t)n_frames, (size_t)S_, s);
                launch_copy_rows(taps_.ratio.p + taps_.fill_vad, ratio_[vbuf_ ^ 1].p, taps_.cap_vad, max_frames_, (size_t)n_frames, (size_t)S_, s);
                FVAD_CUDA_CHECK(cudaMemcpy2DAsync(taps_.met.p + taps_.fill_vad, taps_.cap_vad, met_[vbuf_ ^ 1].p, (size_t)max_frames_,
                                                  (size_t)n_frames, (size_t)S_, cudaMemcpyDeviceToDevice, s));
                taps_.fill_vad += n_frames;
            }
        }
        taps_.fill_chunks += K;
    }

    // ---- carries for the next span ----
    if (want32) launch_copy_rows(ctx_carry_.p, feat_.p + rows * kBins, (size_t)kCtx * kBins, frow * kBins, (size_t)kCtx * kBins, (size_t)SC_, s);
    if (want16) launch_copy_rows(ctx16_carry_.p, feat16_.p + rows * 88, (size_t)kCtx * 88, frow * 88, (size_t)kCtx * 88, (size_t)SC_, s);
    if (full) {
        launch_copy_rows(ylook_carry_.p, y16_.p + ypitch - kYLook, kYLook, ypitch, kYLook, (size_t)SC_, s);
        launch_copy_rows(sumsq_carry_.p, sumsq_.p + K, 1, (size_t)K + 1, 1, (size_t)SC_, s);
    }
    parity_ ^= 1;
    processed_ = span_end;
    frames_done_ += (uint64_t)n_frames;
    chunks_processed_ += (uint64_t)K * SC_;
}

// Queue the held scan on scan_stream_: after its own span's K3b, and after `also_after` when given.
void Batch::issue_scan(cudaEvent_t also_after) {
    const PendingScan p = pscan_;
    pscan_.valid = false;
    cudaEvent_t *ev = span_ev_[p.span].data();
    FVAD_CUDA_CHECK(cudaStreamWaitEvent(scan_stream_, ev[4], 0));
    if (also_after) FVAD_CUDA_CHECK(cudaStreamWaitEvent(scan_stream_, also_after, 0));
    FVAD_CUDA_CHECK(cudaEventRecord(ev[5], scan_stream_));
    if (p.n_frames > 0) {
        ScanArgs ca;
        ca.band = band_[p.buf].p; ca.ratio = ratio_[p.buf].p; ca.machines = d_machines_.p; ca.states = d_states_.p;
        ca.rings = rings_.p; ca.ring_pitch = ring_pitch_; ca.segments = d_segments_.p; ca.max_segments = max_segments_;
        ca.met = met_[p.buf].p; ca.S = S_; ca.M = M_; ca.n_frames = p.n_frames; ca.max_frames = max_frames_;
        ca.first_frame = p.first_frame; ca.fft_size = fft_size_; ca.fast = scan_fast_;
        launch_vadscan(ca, scan_stream_);
    }
    FVAD_CUDA_CHECK(cudaEventRecord(ev[6], scan_stream_));
    FVAD_CUDA_CHECK(cudaEventRecord(scan_done_, scan_stream_));
    if (p.n_frames > 0) FVAD_CUDA_CHECK(cudaEventRecord(scan_done_b_[p.buf], scan_stream_));
}

// Ingest + run.  Host memory (and 16-bit PCM) goes through two stage buffers: the copy stream fills
// buffer b with span i+1 while the compute stream runs span i out of buffer b^1, so H2D overlaps the
// kernels.  Device f32 pointers with nothing pending are used in place.
void Batch::push_any(const void *const *pcm, size_t n_pointers, size_t n, int memory, int fmt) {
    if (n_pointers != (size_t)SC_) throw Error(FVAD_ERR_CHANNEL_COUNT_MISMATCH, "expected n_streams*n_channels planar pointers");
    if (!pcm) throw Error(FVAD_ERR_INVALID_ARGUMENT, "pcm is NULL");
    for (int i = 0; i < SC_; ++i) {
        if (!pcm[i] && n) throw Error(FVAD_ERR_INVALID_ARGUMENT, "NULL channel pointer");
        if (memory == 1 && fmt == 0 && (reinterpret_cast<uintptr_t>(pcm[i]) & 15)) throw Error(FVAD_ERR_INVALID_ARGUMENT, "device PCM pointers must be 16-byte aligned");
    }
    FVAD_CUDA_CHECK(cudaSetDevice(opt_.device));
    total_written_ += n;
    if (skip_ || n == 0) return;  // AudioPipeline.maybeRunPipeline: skip_processing (AudioPipeline.zig:211-214)

    cudaStream_t s = stream_, cs = copy_stream_;
    const cudaMemcpyKind kind = memory == 1 ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    const size_t esz = fmt ? sizeof(int16_t) : sizeof(float);
    const size_t total = (size_t)n_pend_ + n;
    const size_t Ktot = total / (size_t)chunk_;
    const size_t left = total % (size_t)chunk_;
    auto src_at = [&](int i, size_t off) { return static_cast<const char *>(pcm[i]) + off * esz; };

    // taps cover exactly this push
    if (opt_.keep_taps) {
        taps_.fill_chunks = taps_.fill_vad = 0;
        taps_.cap_chunks = Ktot;
        taps_.cap_vad = (size_t)((processed_ + Ktot * chunk_) / (uint64_t)fft_size_ - frames_done_);
        const size_t r = (size_t)SC_ * kFrames * Ktot * kBins;
        taps_.spec.ensure(std::max<size_t>(r, 1)); taps_.feat.ensure(std::max<size_t>(r, 1)); taps_.gains.ensure(std::max<size_t>(r, 1));
        taps_.y16.ensure(std::max<size_t>((size_t)SC_ * kChunk16 * Ktot, 1));
        taps_.den.ensure(std::max<size_t>((size_t)SC_ * chunk_ * Ktot, 1));
        taps_.band.ensure(std::max<size_t>((size_t)S_ * taps_.cap_vad, 1));
        taps_.ratio.ensure(std::max<size_t>((size_t)S_ * taps_.cap_vad, 1));
        taps_.met.ensure(std::max<size_t>((size_t)S_ * taps_.cap_vad, 1));
    }
    if (fmt) pend_i16_.ensure((size_t)SC_ * chunk_);

    if (pscan_.valid) issue_scan(nullptr);  // (only after a push that threw half-way)
    span_idx_ = 0;
    FVAD_CUDA_CHECK(cudaEventRecord(ev_[6], s));
    size_t src_off = 0;        // samples of the caller's buffers consumed by whole chunks
    size_t pend_dst = 0;       // where the leftover goes inside pend_ rows
    if (Ktot > 0 && memory == 1 && fmt == 0 && n_pend_ == 0) {
        // ---- zero-copy: the caller's device buffers are the span input ----
        std::vector<const float *> hx((size_t)SC_);
        for (int i = 0; i < SC_; ++i) hx[(size_t)i] = static_cast<const float *>(pcm[i]);
        FVAD_CUDA_CHECK(cudaMemcpyAsync(d_xptr_.p, hx.data(), hx.size() * sizeof(float *), cudaMemcpyHostToDevice, s));
        FVAD_CUDA_CHECK(cudaStreamSynchronize(s));  // hx is a stack-lifetime pageable buffer
        for (size_t k0 = 0; k0 < Ktot; k0 += (size_t)Kmax_) {
            const int K = (int)std::min<size_t>((size_t)Kmax_, Ktot - k0);
            process_span(d_xptr_.p, k0 * (size_t)chunk_, K);
        }
        src_off = Ktot * (size_t)chunk_;
    } else if (Ktot > 0) {
        // ---- staged, double-buffered ----
        const size_t Kst = std::min<size_t>((size_t)Kmax_, Ktot);
        const size_t spitch = (Kst * (size_t)chunk_ + 7) & ~(size_t)7;
        const int nbuf = Ktot > Kst ? 2 : 1;
        bool moved = spitch != stage_pitch_;
        for (int b = 0; b < nbuf; ++b) {
            const float *old = stage_[b].p;
            stage_[b].ensure((size_t)SC_ * spitch);
            if (fmt) stage_i16_[b].ensure((size_t)SC_ * spitch);
            moved = moved || old != stage_[b].p || !d_sptr_[b].p;
        }
        if (moved) {
            stage_pitch_ = spitch;
            std::vector<const float *> hx((size_t)SC_);
            for (int b = 0; b < nbuf; ++b) {
                d_sptr_[b].ensure((size_t)SC_);
                for (int i = 0; i < SC_; ++i) hx[(size_t)i] = stage_[b].p + (size_t)i * spitch;
                FVAD_CUDA_CHECK(cudaMemcpy(d_sptr_[b].p, hx.data(), hx.size() * sizeof(float *), cudaMemcpyHostToDevice));
            }
        } else if (nbuf == 2 && !d_sptr_[1].p) {
            throw Error(FVAD_ERR_CUDA, "stage pointer table missing");
        }
        // Span schedule: full spans, except a short first one (its copy cannot overlap any compute) and a short
        // last one (its compute cannot overlap any copy) when the push is long enough to have several.
        std::vector<size_t> sched;
        {
            const size_t ramp = std::max<size_t>(Kst / 8, 1);
            size_t rem = Ktot;
            if (Ktot > Kst && !opt_.max_span_chunks) { sched.push_back(ramp); rem -= ramp; }
            while (rem > Kst) { sched.push_back(Kst); rem -= Kst; }
            if (rem > 2 * ramp && Ktot > Kst && !opt_.max_span_chunks) { sched.push_back(rem - ramp); rem = ramp; }
            if (rem) sched.push_back(rem);
        }
        size_t k0 = 0;
        for (size_t span = 0; span < sched.size(); k0 += sched[span], ++span) {
            const int b = (int)(span & 1) % nbuf;
            const size_t K = sched[span];
            const size_t need = K * (size_t)chunk_;
            const size_t from_pend = (k0 == 0) ? (size_t)n_pend_ : 0;
            FVAD_CUDA_CHECK(cudaStreamWaitEvent(cs, stage_free_[b], 0));  // the span that last used buffer b is done
            if (from_pend)
                FVAD_CUDA_CHECK(cudaMemcpy2DAsync(stage_[b].p, spitch * sizeof(float), pend_.p, (size_t)chunk_ * sizeof(float),
                                                  from_pend * sizeof(float), (size_t)SC_, cudaMemcpyDeviceToDevice, cs));
            for (int i = 0; i < SC_; ++i) {
                void *dst = fmt ? static_cast<void *>(stage_i16_[b].p + (size_t)i * spitch)
                                : static_cast<void *>(stage_[b].p + (size_t)i * spitch + from_pend);
                FVAD_CUDA_CHECK(cudaMemcpyAsync(dst, src_at(i, src_off), (need - from_pend) * esz, kind, cs));
            }
            src_off += need - from_pend;
            FVAD_CUDA_CHECK(cudaEventRecord(stage_ready_[b], cs));
            FVAD_CUDA_CHECK(cudaStreamWaitEvent(s, stage_ready_[b], 0));
            if (fmt) launch_i16_to_f32(stage_[b].p + from_pend, stage_i16_[b].p, spitch, spitch, need - from_pend, (size_t)SC_, s);
            process_span(d_sptr_[b].p, 0, (int)K);
            FVAD_CUDA_CHECK(cudaEventRecord(stage_free_[b], s));
        }
    } else {
        pend_dst = (size_t)n_pend_;  // not a whole chunk yet: append
    }
    // ---- the samples that do not fill a chunk wait in pend_ (VADPipeline.zig:150: never processed early) ----
    const size_t n_left = Ktot > 0 ? left : n;
    if (n_left) {
        for (int i = 0; i < SC_; ++i) {
            void *dst = fmt ? static_cast<void *>(pend_i16_.p + (size_t)i * chunk_)
                            : static_cast<void *>(pend_.p + (size_t)i * chunk_ + pend_dst);
            FVAD_CUDA_CHECK(cudaMemcpyAsync(dst, src_at(i, src_off), n_left * esz, kind, cs));
        }
    }
    FVAD_CUDA_CHECK(cudaEventRecord(stage_ready_[0], cs));
    FVAD_CUDA_CHECK(cudaStreamWaitEvent(s, stage_ready_[0], 0));
    if (n_left && fmt) launch_i16_to_f32(pend_.p + pend_dst, pend_i16_.p, (size_t)chunk_, (size_t)chunk_, n_left, (size_t)SC_, s);
    n_pend_ = (int)left;

    if (pscan_.valid) issue_scan(nullptr);
    FVAD_CUDA_CHECK(cudaStreamWaitEvent(s, scan_done_, 0));
    if (Ktot > 0 && !opt_.stft_only) fetch_segments();
    FVAD_CUDA_CHECK(cudaEventRecord(ev_[7], s));
    FVAD_CUDA_CHECK(cudaStreamSynchronize(s));
    float ms = 0;
    cudaEventElapsedTime(&ms, ev_[6], ev_[7]);
    last_ms_[0] = ms;
    if (Ktot > 0) {
        const int last = opt_.stft_only ? 1 : 5;
        for (int i = 0; i < 5; ++i) last_ms_[1 + i] = 0;
        for (size_t sp = 0; sp < span_idx_; ++sp)
            for (int i = 0; i < last; ++i) {
                const int e0 = i < 4 ? i : 5, e1 = i < 4 ? i + 1 : 6;  // the scan runs on scan_stream_
                cudaEventElapsedTime(&ms, span_ev_[sp][e0], span_ev_[sp][e1]);
                last_ms_[1 + i] += ms;
            }
    }
    if (nn_ && Ktot > 0) nn_->take_layer_ms(last_layer_ms_, last_layer_launches_);
    if (segment_overflow_) throw Error(FVAD_ERR_OUT_OF_MEMORY, "segment buffer overflow (more segments than max_segments)");
}

void Batch::fetch_segments() {
    const size_t SM = (size_t)S_ * M_;
    FVAD_CUDA_CHECK(cudaMemcpyAsync(h_states_.data(), d_states_.p, SM * sizeof(ScanState), cudaMemcpyDeviceToHost, stream_));
    FVAD_CUDA_CHECK(cudaStreamSynchronize(stream_));
    // copy only the new segments of each (stream, machine)
    if (seg_seen_.size() != SM) seg_seen_.assign(SM, 0);
    for (size_t t = 0; t < SM; ++t) {
        unsigned have = h_states_[t].n_segments;
        if (have > (unsigned)max_segments_) { segment_overflow_ = true; have = (unsigned)max_segments_; }
        if (have > seg_seen_[t]) {
            FVAD_CUDA_CHECK(cudaMemcpyAsync(h_segments_.data() + t * max_segments_ + seg_seen_[t],
                                            d_segments_.p + t * max_segments_ + seg_seen_[t],
                                            (have - seg_seen_[t]) * sizeof(fvad_segment), cudaMemcpyDeviceToHost, stream_));
            seg_seen_[t] = have;
        }
    }
}

void Batch::segments(size_t stream, size_t machine, const fvad_segment **segs, size_t *n) {
    if (stream >= (size_t)S_ || machine >= (size_t)M_) throw Error(FVAD_ERR_INVALID_ARGUMENT, "stream/machine index out of range");
    const size_t t = stream * M_ + machine;
    if (segs) *segs = h_segments_.data() + t * max_segments_;
    if (n) *n = seg_seen_.empty() ? 0 : seg_seen_[t];
}

size_t Batch::tap(int which, void *dst, size_t dst_bytes) {
    const size_t rows = (size_t)SC_ * kFrames * taps_.fill_chunks * kBins;
    const void *src = nullptr;
    size_t elems = 0, esz = 4;
    switch (which) {
        case FVAD_TAP_SPECTRUM: src = taps_.spec.p; elems = rows; esz = sizeof(fvad_complex); break;
        case FVAD_TAP_FEATURES: src = taps_.feat.p; elems = rows; break;
        case FVAD_TAP_GAINS: src = taps_.gains.p; elems = rows; break;
        case FVAD_TAP_DENOISED16: src = taps_.y16.p; elems = (size_t)SC_ * kChunk16 * taps_.fill_chunks; break;
        case FVAD_TAP_DENOISED: src = taps_.den.p; elems = (size_t)SC_ * chunk_ * taps_.fill_chunks; break;
        case FVAD_TAP_BAND: src = taps_.band.p; elems = (size_t)S_ * taps_.fill_vad; break;
        case FVAD_TAP_RATIO: src = taps_.ratio.p; elems = (size_t)S_ * taps_.fill_vad; break;
        case FVAD_TAP_MET: src = taps_.met.p; elems = (size_t)S_ * taps_.fill_vad; esz = 1; break;
        default: throw Error(FVAD_ERR_INVALID_ARGUMENT, "unknown tap");
    }
    if (!opt_.keep_taps) throw Error(FVAD_ERR_INVALID_ARGUMENT, "taps were not enabled (fvad_batch_options.keep_taps)");
    if (dst) {
        if (dst_bytes < elems * esz) throw Error(FVAD_ERR_INVALID_RESULT_LENGTH, "tap destination too small");
        if (elems) FVAD_CUDA_CHECK(cudaMemcpy(dst, src, elems * esz, cudaMemcpyDeviceToHost));
    }
    return elems;
}

void Batch::stats(fvad_batch_stats *out) const {
    memset(out, 0, sizeof(*out));
    out->kernel_launches = g_kernel_launches - launches0_;
    out->chunks_processed = chunks_processed_;
    out->vad_frames = frames_done_;
    out->last_push_device_ms = last_ms_[0];
    out->last_stft_ms = last_ms_[1]; out->last_nn_ms = last_ms_[2]; out->last_istft_ms = last_ms_[3];
    out->last_vadfft_ms = last_ms_[4]; out->last_scan_ms = last_ms_[5];
    for (int i = 0; i < 9; ++i) { out->last_nn_layer_ms[i] = last_layer_ms_[i]; out->last_nn_layer_launches[i] = last_layer_launches_[i]; }
}

}  // namespace fvad
