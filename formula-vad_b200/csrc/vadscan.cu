// vadscan.cu -- K4: the VAD hysteresis / segment state machine as a per-(stream, machine) scan.
//
// Replaces (reference):
//   RollingAverage.push/avg   (src/structures/RollingAverage.zig:34-56)  f64 means over rings
//   VADMachine.run            (src/AudioPipeline/VADMachine.zig:138-239) thresholds + 4-state FSM
//   VADMachine.trackSpeechStats / onSpeechEnd / getOffsetRecording{Start,End}  (:241-325)
//   VADPipeline.stateMachineStep alt loop (src/AudioPipeline/VADPipeline.zig:231-236): M machines
//   with different Configs over the same frame results (BASELINE config 5).
//
// The reference recomputes each rolling mean as sum_i data[i]*(1/n) on every push (O(window));
// here each ring keeps an f64 running sum (add the new value, subtract the evicted one).  The ring
// values are f32 numbers widened to f64, so the running sum is exact whenever the window's dynamic
// range is below 2^29 and otherwise differs from the reference's own rounding by ~1e-13 relative.
//
// One thread per (stream, machine): the scan is a true recurrence (the long-term mean is frozen
// while the threshold is met), 46.875 steps per stream-second.  Latency-bound by construction;
// the engine overlaps it with the next span's network on another CUDA stream.
#include "kernels.cuh"

namespace fvad {

enum { ST_CLOSED = 0, ST_OPENING = 1, ST_OPEN = 2, ST_CLOSING = 3 };

__device__ __forceinline__ double ring_push(double *ring, int len, int &idx, int &cnt, double &sum, float v) {
    const double nv = (double)v;
    const double old = (cnt == len) ? ring[idx] : 0.0;
    sum = (sum - old) + nv;
    ring[idx] = nv;
    idx = (idx + 1 == len) ? 0 : idx + 1;
    if (cnt < len) ++cnt;
    return sum * (1.0 / (double)cnt);
}

__global__ void __launch_bounds__(128) vadscan_kernel(ScanArgs a) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.S * a.M) return;
    const int s = t / a.M, m = t % a.M;
    const ScanMachine mc = a.machines[m];
    ScanState st = a.states[t];
    double *ring_long = a.rings + (size_t)t * a.ring_pitch;
    double *ring_short = ring_long + mc.n_long;
    double *ring_ratio = ring_short + mc.n_short;
    const float *__restrict__ band = a.band + ((size_t)mc.band * a.S + s) * a.max_frames;
    const float *__restrict__ ratio = a.ratio + (size_t)s * a.max_frames;
    fvad_segment *segs = a.segments + (size_t)t * a.max_segments;
    unsigned char *met_out = (a.met && m == 0) ? a.met + (size_t)s * a.max_frames : nullptr;

    constexpr int PF = 8;  // frames fetched ahead of the dependent chain
    for (int i0 = 0; i0 < a.n_frames; i0 += PF) {
        float eb[PF], rb[PF];
#pragma unroll
        for (int j = 0; j < PF; ++j) {
            const int i = min(i0 + j, a.n_frames - 1);
            eb[j] = band[i];
            rb[j] = ratio[i];
        }
#pragma unroll
        for (int j = 0; j < PF; ++j) {
            const int i = i0 + j;
            if (i >= a.n_frames) break;
            const float min_volume = eb[j], meta_ratio = rb[j];
            const unsigned long long index = (unsigned long long)(a.first_frame + i) * (unsigned long long)a.fft_size;

            // VADMachine.zig:165-178
            const double short_term = ring_push(ring_short, mc.n_short, st.short_idx, st.short_cnt, st.short_sum, min_volume);
            const double vol_ratio = ring_push(ring_ratio, mc.n_ratio, st.ratio_idx, st.ratio_cnt, st.ratio_sum, meta_ratio);
            const double base = st.has_last ? st.last_avg : (mc.has_init ? mc.init_avg : short_term);
            const double threshold = base * mc.threshold_factor;
            const bool met = (short_term > threshold) && (vol_ratio > mc.ratio_threshold);
            if (!met) {
                st.last_avg = ring_push(ring_long, mc.n_long, st.long_idx, st.long_cnt, st.long_sum, min_volume);
                st.has_last = 1;
            }
            if (met_out) met_out[i] = met ? 1 : 0;

            const int from_state = st.state;
            switch (st.state) {  // VADMachine.zig:189-233
                case ST_CLOSED:
                    if (met) { st.state = ST_OPENING; st.start_index = index; }
                    break;
                case ST_OPENING:
                    if (met && (index - st.start_index) >= mc.min_to_open) st.state = ST_OPEN;
                    else if (!met) st.state = ST_CLOSED;
                    break;
                case ST_OPEN:
                    if (!met) { st.state = ST_CLOSING; st.end_index = index; }
                    break;
                case ST_CLOSING:
                    if (met) st.state = ST_OPEN;
                    else if ((index - st.end_index) >= mc.max_gap) {
                        st.state = ST_CLOSED;
                        // onSpeechEnd, VADMachine.zig:265-309
                        const float len_sec = __fdiv_rn((float)(st.end_index - st.start_index), mc.sample_rate_f);
                        if (len_sec >= mc.min_duration_sec) {
                            if (st.n_segments < (unsigned)a.max_segments) {
                                fvad_segment g;
                                const unsigned long long mg = mc.start_margin < st.start_index ? mc.start_margin : st.start_index;
                                g.sample_from = st.start_index - mg;
                                g.sample_to = st.end_index + mc.end_margin;
                                g.avg_channel_vol_ratio = __fdiv_rn(st.vol_ratio_sum, (float)st.vol_ratio_count);
                                g.vad_met_sec = st.met_sec;
                                segs[st.n_segments] = g;
                            }
                            ++st.n_segments;
                        }
                    }
                    break;
            }
            // trackSpeechStats, VADMachine.zig:241-263
            if (from_state == ST_CLOSED && st.state == ST_OPENING) {
                st.vol_ratio_sum = meta_ratio;
                st.vol_ratio_count = 1;
                st.met_sec = mc.frame_sec;
            } else if (from_state == ST_OPEN) {
                st.vol_ratio_sum = __fadd_rn(st.vol_ratio_sum, meta_ratio);
                st.vol_ratio_count += 1;
                if (met) st.met_sec = __fadd_rn(st.met_sec, mc.frame_sec);
            }
        }
    }
    a.states[t] = st;
}

void launch_vadscan(const ScanArgs &a, cudaStream_t s) {
    if (a.n_frames <= 0) return;
    const int total = a.S * a.M;
    vadscan_kernel<<<(total + 127) / 128, 128, 0, s>>>(a);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace fvad
