// vadscan.cu -- K4: the VAD hysteresis / segment state machine as a per-(stream, machine) scan.
//
// Replaces (reference):
//   RollingAverage.push/avg   (src/structures/RollingAverage.zig:34-56)  f64 means over rings
//   VADMachine.run            (src/AudioPipeline/VADMachine.zig:138-239) thresholds + 4-state FSM
//   VADMachine.trackSpeechStats / onSpeechEnd / getOffsetRecording{Start,End}  (:241-325)
//   VADPipeline.stateMachineStep alt loop (src/AudioPipeline/VADPipeline.zig:231-236): M machines
//   with different Configs over the same frame results (BASELINE config 5).
//
// The reference recomputes each rolling mean as sum_i data[i]*(1/n) on every push (O(window)).
// Here the short-term and ratio means are evaluated exactly so (fast path); the long-term ring
// (8 437 slots) keeps an f64 running sum -- add the new value, subtract the evicted one -- that is
// RE-DERIVED from the ring every time the write index wraps (once per n_long pushes, ~3 min of
// audio), so rounding never accumulates beyond one turn of the ring.  The ring values are f32
// numbers widened to f64: the running sum is exact while the window's dynamic range stays below
// 2^29 / n_long (then it equals the reference's sum up to the reference's own per-term rounding)
// and otherwise carries at most one window's worth of f64 rounding.
//
// Two kernels: `vadscan_warp_kernel` (the engine's path) runs one WARP per (stream, machine) and
// advances the long-term recurrence by runs of frames; `vadscan_kernel` (window sizes outside the
// fast path's limits) runs one thread per (stream, machine), frame by frame.  The scan is a true
// recurrence (the long-term mean is frozen while the threshold is met), 46.875 steps per
// stream-second: latency-bound by construction; the engine overlaps it with the next span's
// network on another CUDA stream.
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

// FSM + statistics of one frame, given this frame's threshold decision (VADMachine.zig:189-263)
__device__ __forceinline__ void fsm_step(ScanState &st, const ScanMachine &mc, bool met, float meta_ratio,
                                         unsigned long long index, fvad_segment *segs, int max_segments) {
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
                    if (st.n_segments < (unsigned)max_segments) {
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

constexpr int kScanThreads = 32;
constexpr int kSmallRing = 64;   // doubles of shared memory for the short + ratio rings (fast path)
constexpr int kBatch = 256;      // frames staged per batch (fast path)
constexpr int PF = 8;            // frames fetched ahead of the dependent chain (generic path)

// Generic path: one thread per (stream, machine), rings in global memory.
__global__ void __launch_bounds__(kScanThreads) vadscan_kernel(ScanArgs a) {
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
    unsigned char *met_out = !a.met ? nullptr : a.met_per_machine ? a.met + (size_t)t * a.max_frames
                             : (m == 0 ? a.met + (size_t)s * a.max_frames : nullptr);

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
                if (st.long_idx == 0) {  // the ring has turned over once: re-derive the running sum from its contents
                    double acc = 0.0;
                    for (int k = 0; k < mc.n_long; ++k) acc += ring_long[k];
                    st.long_sum = acc;
                    st.last_avg = acc * (1.0 / (double)st.long_cnt);
                }
            }
            if (met_out) met_out[i] = met ? 1 : 0;
            fsm_step(st, mc, met, meta_ratio, index, segs, a.max_segments);
        }
    }
    a.states[t] = st;
}

// Fast path: one WARP per (stream, machine), frames staged through shared memory in batches.
//   * The short-term and ratio means do not depend on the recurrence: all lanes evaluate them in
//     parallel, each as the reference does -- the array-order sum of data[i] * (1/written)
//     (RollingAverage.zig:45-56) -- from a chronological history kept in shared memory, so these two
//     means are bit-identical to the reference's.  (In this mode the short/ratio ring storage holds
//     the last n-1 pushed values in chronological order instead of a circular buffer.)
//   * The long-term mean is the true recurrence (it is only pushed on !met frames): lane 0 runs it
//     as an f64 running sum; the next kBatch+2 eviction candidates of the long ring are staged
//     ahead (the ring advances by at most one slot per frame and n_long >= 2*kBatch+2), so no
//     global-memory latency sits on the chain last_avg -> threshold -> met -> long sum -> last_avg.
__device__ __forceinline__ double window_mean(const float *newest, int n, int pm, int cnt) {
    // data[] position i holds the value pushed r = (pm - i) mod n pushes ago; sum in array order
    const double scalar = 1.0 / (double)cnt;
    double acc = 0.0;
    for (int i = 0; i < cnt; ++i) {
        int r = pm - i;
        if (r < 0) r += n;
        acc = __dadd_rn(acc, __dmul_rn((double)newest[-r], scalar));
    }
    return acc;
}

// The long-term recurrence, run by RUNS instead of by frames.  Between two threshold crossings the recurrence
// has a closed form: while no frame meets the threshold every frame is pushed, so the running sum before frame
// l is sum + prefix(new - evicted) and all 32 thresholds of a block come from one warp prefix sum; while every
// frame meets it nothing changes at all.  Each iteration assumes "no frame met" from the cursor on, finds the
// first lane that contradicts it, commits the pushes before that lane, then assumes "every frame met" and finds
// the first lane that contradicts that.  Speech and silence come in runs of hundreds of frames, so a block of
// 32 frames usually takes one or two iterations (~10 cycles a frame against ~95 for the frame-by-frame chain);
// after kMaxIters iterations in one block (a decision flickering frame by frame) lane 0 finishes the block
// sequentially.  The prefix sum associates the f64 additions differently from the sequential running sum: both
// are exact while the window's dynamic range stays below 2^29 / n_long and differ by ~1e-16 relative beyond.
constexpr int kMaxIters = 6;

struct LongRun {       // warp-uniform copy of the long-term ring's state
    double sum, last_avg, inv;
    int idx, cnt, has_last;
    int k;             // pushes since the eviction candidates were staged
};

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

__global__ void __launch_bounds__(kScanThreads) vadscan_warp_kernel(ScanArgs a) {
    __shared__ float s_rb[kBatch];
    __shared__ double s_lw[kBatch + 2], s_st[kBatch], s_vr[kBatch];
    __shared__ float x_s[kSmallRing + kBatch], x_r[kSmallRing + kBatch];  // [n-1 history | batch], chronological
    const int t = blockIdx.x, lane = threadIdx.x;
    const int s = t / a.M, m = t % a.M;
    const ScanMachine mc = a.machines[m];
    ScanState st = a.states[t];
    double *ring_long = a.rings + (size_t)t * a.ring_pitch;
    double *ring_short = ring_long + mc.n_long;
    double *ring_ratio = ring_short + mc.n_short;
    const float *__restrict__ band = a.band + ((size_t)mc.band * a.S + s) * a.max_frames;
    const float *__restrict__ ratio = a.ratio + (size_t)s * a.max_frames;
    fvad_segment *segs = a.segments + (size_t)t * a.max_segments;
    unsigned char *met_out = !a.met ? nullptr : a.met_per_machine ? a.met + (size_t)t * a.max_frames
                             : (m == 0 ? a.met + (size_t)s * a.max_frames : nullptr);
    const int hs = mc.n_short - 1, hr = mc.n_ratio - 1;
    for (int k = lane; k < hs; k += 32) x_s[k] = (float)ring_short[k];
    for (int k = lane; k < hr; k += 32) x_r[k] = (float)ring_ratio[k];
    unsigned long long index = (unsigned long long)a.first_frame * (unsigned long long)a.fft_size;
    LongRun L;
    L.sum = st.long_sum; L.last_avg = st.last_avg; L.idx = st.long_idx; L.cnt = st.long_cnt; L.has_last = st.has_last;
    L.inv = 1.0 / (double)(L.cnt > 0 ? L.cnt : 1);
    L.k = 0;

    for (int i0 = 0; i0 < a.n_frames; i0 += kBatch) {
        const int nb = min(kBatch, a.n_frames - i0);
        for (int j = lane; j < nb; j += 32) {
            x_s[hs + j] = band[i0 + j];
            const float r = ratio[i0 + j];
            x_r[hr + j] = r;
            s_rb[j] = r;
        }
        for (int j = lane; j < nb + 2; j += 32) {
            int li = L.idx + j;
            if (li >= mc.n_long) li -= mc.n_long;
            s_lw[j] = ring_long[li];
        }
        L.k = 0;
        __syncwarp();
        for (int j = lane; j < nb; j += 32) {
            s_st[j] = window_mean(x_s + hs + j, mc.n_short, (st.short_idx + j) % mc.n_short, min(st.short_cnt + j + 1, mc.n_short));
            s_vr[j] = window_mean(x_r + hr + j, mc.n_ratio, (st.ratio_idx + j) % mc.n_ratio, min(st.ratio_cnt + j + 1, mc.n_ratio));
        }
        __syncwarp();

        for (int jb = 0; jb < nb; jb += 32) {
            const int nl = min(32, nb - jb);
            const int j = jb + lane;
            const bool valid = lane < nl;
            const float nvf = valid ? x_s[hs + j] : 0.0f;
            const double stv = valid ? s_st[j] : 0.0;
            const bool vrok = valid && (s_vr[j] > mc.ratio_threshold);
            unsigned metmask = 0;
            int cur = 0;
            int wrapped = 0;  // the long ring's write index passed slot 0 in this block
            if (L.has_last && L.cnt == mc.n_long) {
                for (int it = 0; it < kMaxIters && cur < nl; ++it) {
                    // ---- no frame from the cursor on meets the threshold: every one of them is pushed ----
                    const int rel = lane - cur;
                    const bool act = rel >= 0 && valid;
                    const double d = act ? (double)nvf - s_lw[L.k + rel] : 0.0;
                    double inc = d;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const double up = __shfl_up_sync(0xffffffffu, inc, o);
                        if (lane >= o) inc += up;
                    }
                    double exc = __shfl_up_sync(0xffffffffu, inc, 1);
                    if (lane == 0) exc = 0.0;
                    const double base = (rel == 0) ? L.last_avg : (L.sum + exc) * L.inv;
                    const bool mA = act && (stv > base * mc.threshold_factor) && vrok;
                    const unsigned hit = __ballot_sync(0xffffffffu, mA);
                    const int f = hit ? __ffs((int)hit) - 1 : nl;
                    const int np = f - cur;
                    if (np > 0) {
                        if (act && lane < f) {
                            int li = L.idx + rel;
                            if (li >= mc.n_long) li -= mc.n_long;
                            ring_long[li] = (double)nvf;
                        }
                        L.sum += shfl_d(inc, f - 1);
                        L.last_avg = L.sum * L.inv;
                        L.idx += np;
                        if (L.idx >= mc.n_long) { L.idx -= mc.n_long; wrapped = 1; }
                        L.k += np;
                    }
                    cur = f;
                    if (cur >= nl) break;
                    // ---- every frame from the cursor on meets it: the long-term mean is frozen ----
                    const bool mB = (stv > L.last_avg * mc.threshold_factor) && vrok;
                    const unsigned stop = __ballot_sync(0xffffffffu, !mB && lane >= cur && valid);
                    const int g = stop ? __ffs((int)stop) - 1 : nl;
                    metmask |= (g >= 32 ? 0xffffffffu : ((1u << g) - 1u)) & ~((1u << cur) - 1u);
                    cur = g;
                }
            }
            if (cur < nl) {
                // frame by frame on lane 0: before the ring is full / the first push, or a flickering block
                if (lane == 0) {
                    for (int l = cur; l < nl; ++l) {
                        const float min_volume = x_s[hs + jb + l];
                        const double short_term = s_st[jb + l];
                        // VADMachine.zig:165-178
                        const double base = L.has_last ? L.last_avg : (mc.has_init ? mc.init_avg : short_term);
                        const bool met = (short_term > base * mc.threshold_factor) && (s_vr[jb + l] > mc.ratio_threshold);
                        if (!met) {
                            const double nv = (double)min_volume;
                            const double old = (L.cnt == mc.n_long) ? s_lw[L.k] : 0.0;
                            L.sum = (L.sum - old) + nv;
                            ring_long[L.idx] = nv;
                            L.idx = (L.idx + 1 == mc.n_long) ? 0 : L.idx + 1;
                            if (L.idx == 0) wrapped = 1;
                            if (L.cnt < mc.n_long) {
                                ++L.cnt;
                                L.inv = 1.0 / (double)L.cnt;
                            }
                            L.last_avg = L.sum * L.inv;
                            L.has_last = 1;
                            ++L.k;
                        } else {
                            metmask |= 1u << l;
                        }
                    }
                }
                L.sum = shfl_d(L.sum, 0); L.last_avg = shfl_d(L.last_avg, 0); L.inv = shfl_d(L.inv, 0);
                L.idx = __shfl_sync(0xffffffffu, L.idx, 0); L.cnt = __shfl_sync(0xffffffffu, L.cnt, 0);
                L.has_last = __shfl_sync(0xffffffffu, L.has_last, 0); L.k = __shfl_sync(0xffffffffu, L.k, 0);
                metmask = __shfl_sync(0xffffffffu, metmask, 0);
                wrapped = __shfl_sync(0xffffffffu, wrapped, 0);
            }
            if (wrapped) {
                // One turn of the ring since the last re-derivation: replace the running sum (n_long add/subtract
                // pairs of rounding) by the sum of what the ring holds now -- what the reference sums on every push
                // (RollingAverage.zig:45-56).  All values are >= 0, so the lane-strided tree sum is within a few ulp
                // of the array-order sum; ~3 us per 3 minutes of audio.
                __syncwarp();  // this block's ring stores (other lanes) are visible
                double part = 0.0;
                for (int k = lane; k < mc.n_long; k += 32) part += ring_long[k];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
                L.sum = part;
                if (L.has_last) L.last_avg = L.sum * L.inv;
            }
            if (met_out && valid) met_out[i0 + j] = (unsigned char)((metmask >> lane) & 1u);
            // ---- state machine + statistics: sequential, but idle while closed with nothing met ----
            if (lane == 0) {
                if (metmask == 0 && st.state == ST_CLOSED) {  // VADMachine.zig:190-196: nothing changes
                    index += (unsigned long long)nl * (unsigned long long)a.fft_size;
                } else {
                    for (int l = 0; l < nl; ++l) {
                        const bool met = (metmask >> l) & 1u;
                        if (met || st.state != ST_CLOSED) fsm_step(st, mc, met, s_rb[jb + l], index, segs, a.max_segments);
                        index += (unsigned long long)a.fft_size;
                    }
                }
            }
        }
        // ring counters advance identically on every lane; histories shift by nb
        st.short_idx = (st.short_idx + nb) % mc.n_short;
        st.short_cnt = min(st.short_cnt + nb, mc.n_short);
        st.ratio_idx = (st.ratio_idx + nb) % mc.n_ratio;
        st.ratio_cnt = min(st.ratio_cnt + nb, mc.n_ratio);
        __syncwarp();
        float v0 = 0.f, v1 = 0.f, w0 = 0.f, w1 = 0.f;
        if (lane < hs) v0 = x_s[nb + lane];
        if (lane + 32 < hs) v1 = x_s[nb + lane + 32];
        if (lane < hr) w0 = x_r[nb + lane];
        if (lane + 32 < hr) w1 = x_r[nb + lane + 32];
        __syncwarp();
        if (lane < hs) x_s[lane] = v0;
        if (lane + 32 < hs) x_s[lane + 32] = v1;
        if (lane < hr) x_r[lane] = w0;
        if (lane + 32 < hr) x_r[lane + 32] = w1;
        __syncwarp();
    }
    for (int k = lane; k < hs; k += 32) ring_short[k] = (double)x_s[k];
    for (int k = lane; k < hr; k += 32) ring_ratio[k] = (double)x_r[k];
    if (lane == 0) {
        st.long_sum = L.sum; st.last_avg = L.last_avg; st.long_idx = L.idx; st.long_cnt = L.cnt; st.has_last = L.has_last;
        a.states[t] = st;
    }
}

void launch_vadscan(const ScanArgs &a, cudaStream_t s) {
    if (a.n_frames <= 0) return;
    const int total = a.S * a.M;
    static bool attr = false;
    if (!attr) {
        // The scan runs on its own stream next to the persistent NSNet2 kernels of the following span (~200 KB of
        // shared memory per SM): keep the shared-memory carveout at its maximum so both kinds of CTA fit one SM.
        // (What the scan does disturb is the GRU's issue-bound epilogue; the engine schedules around that.)
        cudaFuncSetAttribute(vadscan_warp_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(vadscan_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        attr = true;
    }
    if (a.fast) vadscan_warp_kernel<<<(unsigned)total, kScanThreads, 0, s>>>(a);
    else vadscan_kernel<<<(unsigned)((total + kScanThreads - 1) / kScanThreads), kScanThreads, 0, s>>>(a);
    FVAD_LAUNCHED();
    FVAD_CUDA_CHECK(cudaGetLastError());
}

}  // namespace fvad
