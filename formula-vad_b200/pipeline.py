"""Host-side mirror of the reference's operator/plugin interface for the hot path, over the C ABI.

Same names, argument meaning and error behaviour as the reference (Zig) so the parity tests read
like the reference's own tests:
    FFT            <- src/FFT.zig            (init / fft / invFft / binCount / binWidth / freqToBin / binToFreq)
    NSNet2         <- src/NSNet2.zig         (init / getChunkSize / denoise / calcSpectrogram)
    AudioPipeline  <- src/AudioPipeline.zig  (init / pushSamples / totalWriteCount / vad_segments)
    BatchPipeline  <- many AudioPipelines advanced in lock step (simulator.zig:221-232 folded into the batch)
Zig error names become FvadError.name (e.g. "InvalidSampleRate").
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .capi import FvadError, check  # noqa: F401


def vad_machine_config(**kw) -> capi.VadMachineConfig:
    """VADMachine.Config with the reference defaults (VADMachine.zig:30-51), overridable by name."""
    c = capi.VadMachineConfig()
    capi.lib().fvad_vad_machine_config_default(C.byref(c))
    for k, v in kw.items():
        if not hasattr(c, k):
            raise AttributeError(k)
        setattr(c, k, v)
    return c


def _config(sample_rate, n_channels, vad_machine_config_=None, alt_vad_machine_configs=None,
            denoiser_model_path=None, skip_processing=False, fft_size=1024):
    cfg = capi.Config()
    capi.lib().fvad_config_default(C.byref(cfg), sample_rate, n_channels)
    cfg.vad_config.fft_size = fft_size
    if vad_machine_config_ is not None:
        cfg.vad_config.vad_machine_config = vad_machine_config_
    keep = []
    if alt_vad_machine_configs:
        arr = (capi.VadMachineConfig * len(alt_vad_machine_configs))(*alt_vad_machine_configs)
        cfg.vad_config.alt_vad_machine_configs = arr
        cfg.vad_config.n_alt_vad_machine_configs = len(alt_vad_machine_configs)
        keep.append(arr)
    if denoiser_model_path is not None:
        b = denoiser_model_path.encode()
        cfg.vad_config.denoiser_model_path = b
        keep.append(b)
    cfg.skip_processing = int(skip_processing)
    return cfg, keep


def _segments_list(ptr, n):
    return [(ptr[i].sample_from, ptr[i].sample_to, ptr[i].avg_channel_vol_ratio, ptr[i].vad_met_sec) for i in range(n)]


class FFT:
    """src/FFT.zig.  n_fft in {64,128,256,320,512,1024} on the GPU."""
    _h = None

    def __init__(self, n_fft: int, sample_rate: int, mode_inverse: bool = False):
        self._L = capi.lib()
        h = C.c_void_p()
        check(self._L.fvad_fft_create(n_fft, sample_rate, int(mode_inverse), C.byref(h)))
        self._h, self.n_fft, self.sample_rate = h, n_fft, sample_rate

    def fft(self, first: np.ndarray, window: np.ndarray, second: np.ndarray | None = None, n_bins: int | None = None):
        """FFT.fft(SplitSlice{first, second}, window, bins)."""
        first = np.ascontiguousarray(first, np.float32)
        second = np.zeros(0, np.float32) if second is None else np.ascontiguousarray(second, np.float32)
        window = np.ascontiguousarray(window, np.float32)
        nb = self.bin_count() if n_bins is None else n_bins
        bins = np.empty(nb, np.complex64)
        check(self._L.fvad_fft_forward(self._h, first.ctypes.data_as(capi.c_f32p), first.size,
                                       second.ctypes.data_as(capi.c_f32p), second.size,
                                       window.ctypes.data_as(capi.c_f32p), window.size,
                                       bins.ctypes.data_as(C.POINTER(capi.Complex)), nb))
        return bins

    def inv_fft(self, bins: np.ndarray, n_result: int | None = None):
        bins = np.ascontiguousarray(bins, np.complex64)
        n = self.n_fft if n_result is None else n_result
        out = np.empty(n, np.float32)
        check(self._L.fvad_fft_inverse(self._h, bins.ctypes.data_as(C.POINTER(capi.Complex)), bins.size,
                                       out.ctypes.data_as(capi.c_f32p), n))
        return out

    def fft_batch(self, frames: np.ndarray, window: np.ndarray | None = None):
        frames = np.ascontiguousarray(frames, np.float32)
        nf = frames.shape[0]
        bins = np.empty((nf, self.bin_count()), np.complex64)
        w = None if window is None else np.ascontiguousarray(window, np.float32)
        check(self._L.fvad_fft_forward_batch(self._h, frames.ctypes.data, nf, frames.shape[1],
                                             None if w is None else w.ctypes.data, bins.ctypes.data, 0))
        return bins

    def inv_fft_batch(self, bins: np.ndarray):
        bins = np.ascontiguousarray(bins, np.complex64)
        out = np.empty((bins.shape[0], self.n_fft), np.float32)
        check(self._L.fvad_fft_inverse_batch(self._h, bins.ctypes.data, bins.shape[0], out.ctypes.data, 0))
        return out

    def bin_count(self) -> int:
        return self._L.fvad_fft_bin_count(self._h)

    def bin_width(self) -> float:
        return self._L.fvad_fft_bin_width(self._h)

    def nyquist_freq(self) -> float:
        return self._L.fvad_fft_nyquist_freq(self._h)

    def freq_to_bin(self, freq: float) -> int:
        b = C.c_size_t()
        check(self._L.fvad_fft_freq_to_bin(self._h, freq, C.byref(b)))
        return b.value

    def bin_to_freq(self, b: int) -> float:
        f = C.c_float()
        check(self._L.fvad_fft_bin_to_freq(self._h, b, C.byref(f)))
        return f.value

    def deinit(self):
        if self._h:
            self._L.fvad_fft_destroy(self._h)
            self._h = None

    __del__ = deinit


class NSNet2:
    """src/NSNet2.zig: one denoiser per channel, one chunk per call, stateful."""
    _h = None

    def __init__(self, sample_rate: int, model_path: str | None = None, weights: dict | None = None,
                 nn_mode: int = capi.NN_F32):
        self._L = capi.lib()
        h = C.c_void_p()
        if weights is not None:
            w, self._keep = capi.make_weights(weights)
            check(self._L.fvad_nsnet2_create_with_weights(sample_rate, C.byref(w), nn_mode, C.byref(h)))
        else:
            check(self._L.fvad_nsnet2_create(sample_rate, None if model_path is None else model_path.encode(), C.byref(h)))
        self._h, self.sample_rate = h, sample_rate

    @staticmethod
    def get_chunk_size(sample_rate: int) -> int:
        return capi.lib().fvad_nsnet2_chunk_size(sample_rate)

    def denoise(self, first: np.ndarray, second: np.ndarray | None = None, n_result: int | None = None):
        first = np.ascontiguousarray(first, np.float32)
        second = np.zeros(0, np.float32) if second is None else np.ascontiguousarray(second, np.float32)
        n = first.size + second.size if n_result is None else n_result
        out = np.empty(n, np.float32)
        check(self._L.fvad_nsnet2_denoise(self._h, first.ctypes.data_as(capi.c_f32p), first.size,
                                          second.ctypes.data_as(capi.c_f32p), second.size,
                                          out.ctypes.data_as(capi.c_f32p), n))
        return out

    @staticmethod
    def calc_spectrogram(audio_chunk: np.ndarray, n_frames: int, window: np.ndarray):
        a = np.ascontiguousarray(audio_chunk, np.float32)
        w = np.ascontiguousarray(window, np.float32)
        out = np.empty((n_frames, 161), np.complex64)
        check(capi.lib().fvad_nsnet2_calc_spectrogram(a.ctypes.data_as(capi.c_f32p), a.size, n_frames,
                                                      w.ctypes.data_as(capi.c_f32p), w.size,
                                                      out.ctypes.data_as(C.POINTER(capi.Complex)), out.size))
        return out

    def deinit(self):
        if self._h:
            self._L.fvad_nsnet2_destroy(self._h)
            self._h = None

    __del__ = deinit


class AudioPipeline:
    """src/AudioPipeline.zig facade: init(Config) / pushSamples / totalWriteCount / vad_segments."""
    _h = None

    def __init__(self, sample_rate: int, n_channels: int, weights: dict | None = None, vad_machine_config=None,
                 alt_vad_machine_configs=None, denoiser_model_path: str | None = None, skip_processing=False,
                 nn_mode: int = capi.NN_F32, fft_size: int = 1024, on_original_recording=None,
                 on_denoised_recording=None):
        """on_*_recording(start_index, pcm[channels][n], sample_rate): AudioPipeline.Callbacks (AudioPipeline.zig:14-18);
        the arrays are copies (the C buffers are only borrowed for the call)."""
        self._L = capi.lib()
        self.cfg, self._keep = _config(sample_rate, n_channels, vad_machine_config, alt_vad_machine_configs,
                                       denoiser_model_path, skip_processing, fft_size)
        self.n_channels = n_channels
        cbp = None
        if on_original_recording or on_denoised_recording:
            def wrap(fn):
                def thunk(_ctx, rec):
                    r = rec.contents
                    pcm = np.stack([np.ctypeslib.as_array(r.channel_pcm[c], shape=(r.length,)).copy() for c in range(r.n_channels)])
                    fn(int(r.global_start_frame_number), pcm, int(r.sample_rate))
                return capi.RECORDING_CB(thunk) if fn else capi.RECORDING_CB()
            cbs = capi.Callbacks(None, wrap(on_original_recording), wrap(on_denoised_recording))
            self._keep.append(cbs)
            cbp = C.byref(cbs)
        h = C.c_void_p()
        if weights is not None:
            w, k = capi.make_weights(weights)
            self._keep.append(k)
            check(self._L.fvad_pipeline_create_with_weights(C.byref(self.cfg), cbp, C.byref(w), nn_mode, C.byref(h)))
        else:
            check(self._L.fvad_pipeline_create(C.byref(self.cfg), cbp, C.byref(h)))
        self._h = h

    def push_samples(self, channel_pcm) -> int:
        """channel_pcm: sequence of n_channels float32 arrays (planar).  Returns the absolute index
        of the first sample pushed (AudioPipeline.zig:118-119)."""
        chans = [np.ascontiguousarray(c, np.float32) for c in channel_pcm]
        n = chans[0].size if chans else 0
        ptrs = (capi.c_f32p * len(chans))(*[c.ctypes.data_as(capi.c_f32p) for c in chans])
        first = C.c_uint64()
        check(self._L.fvad_pipeline_push_samples(self._h, ptrs, len(chans), n, C.byref(first)))
        return first.value

    def total_write_count(self) -> int:
        return self._L.fvad_pipeline_total_write_count(self._h)

    def vad_segments(self, machine: int = 0):
        p, n = C.POINTER(capi.Segment)(), C.c_size_t()
        check(self._L.fvad_pipeline_segments(self._h, machine, C.byref(p), C.byref(n)))
        return _segments_list(p, n.value)

    def deinit(self):
        if self._h:
            self._L.fvad_pipeline_destroy(self._h)
            self._h = None

    __del__ = deinit


def eval_config(**kw) -> capi.EvalConfig:
    """StatConfig defaults of the reference (Evaluator/statistics.zig:77-83, beta 0.7) with overrides."""
    c = capi.EvalConfig()
    capi.lib().fvad_eval_config_default(C.byref(c))
    for k, v in kw.items():
        if not hasattr(c, k):
            raise AttributeError(k)
        setattr(c, k, v)
    return c


def _eval_segs(segs):
    arr = (capi.EvalSegment * max(len(segs), 1))()
    for i, (a, b) in enumerate(segs):
        arr[i].from_sec, arr[i].to_sec = a, b
    return arr


def _stats_dict(st) -> dict:
    return {k: getattr(st, k) for k, _ in capi.EvalStats._fields_}


class Evaluator:
    """Mirror of src/Evaluator.zig + Evaluator/statistics.zig: `Evaluator.init_and_run(vad, ref).stats(cfg)`.
    Segments are (from_sec, to_sec) pairs."""

    def __init__(self, input_segments, reference_segments):
        self.input_segments, self.reference_segments = list(input_segments), list(reference_segments)

    @classmethod
    def init_and_run(cls, input_segments, reference_segments):
        return cls(input_segments, reference_segments)

    def stats(self, config: capi.EvalConfig | None = None) -> dict:
        st = capi.EvalStats()
        cfg = config if config is not None else eval_config()
        check(capi.lib().fvad_evaluate(_eval_segs(self.input_segments), len(self.input_segments),
                                       _eval_segs(self.reference_segments), len(self.reference_segments),
                                       C.byref(cfg), C.byref(st)))
        return _stats_dict(st)


class VADMachine:
    """Mirror of src/AudioPipeline/VADMachine.zig run over a whole series: `VADMachine(configs, sample_rate).scan(
    band_min, ratio)`.  band_min is [n_machines][n_frames] (each machine's speech-band sum minimised over the
    channels), ratio [n_frames].  Returns (met [n_machines][n_frames] uint8, [segments per machine])."""

    def __init__(self, configs, sample_rate: int, fft_size: int = 1024, max_segments: int = 4096):
        self.configs = list(configs) if isinstance(configs, (list, tuple)) else [configs]
        self.sample_rate, self.fft_size, self.max_segments = sample_rate, fft_size, max_segments

    def scan(self, band_min: np.ndarray, ratio: np.ndarray, frames_per_launch: int = 0):
        M = len(self.configs)
        band = np.ascontiguousarray(band_min, dtype=np.float32).reshape(M, -1)
        rat = np.ascontiguousarray(ratio, dtype=np.float32).reshape(-1)
        n = band.shape[1]
        if rat.shape[0] != n:
            raise ValueError("ratio must have one value per frame")
        cfgs = (capi.VadMachineConfig * M)(*self.configs)
        met = np.zeros((M, n), dtype=np.uint8)
        segs = (capi.Segment * (M * self.max_segments))()
        nseg = (C.c_size_t * M)()
        check(capi.lib().fvad_vad_machine_scan(cfgs, M, self.sample_rate, self.fft_size, band.ctypes.data_as(capi.c_f32p),
                                               rat.ctypes.data_as(capi.c_f32p), n,
                                               frames_per_launch, met.ctypes.data_as(C.c_void_p), segs, self.max_segments, nseg))
        out = []
        for m in range(M):
            k = min(int(nseg[m]), self.max_segments)
            out.append([(segs[m * self.max_segments + i].sample_from, segs[m * self.max_segments + i].sample_to,
                         segs[m * self.max_segments + i].avg_channel_vol_ratio, segs[m * self.max_segments + i].vad_met_sec)
                        for i in range(k)])
        return met, out


class BatchPipeline:
    """n_streams independent AudioPipelines advanced in lock step on one GPU."""
    _h = None

    def __init__(self, n_streams: int, sample_rate: int, n_channels: int, weights: dict | None = None,
                 vad_machine_config=None, alt_vad_machine_configs=None, nn_mode: int = capi.NN_F32, device: int = 0,
                 keep_taps=False, stft_only=False, max_span_chunks: int = 0, forced_gains: np.ndarray | None = None):
        self._L = capi.lib()
        self.cfg, self._keep = _config(sample_rate, n_channels, vad_machine_config, alt_vad_machine_configs)
        self.n_streams, self.n_channels, self.sample_rate = n_streams, n_channels, sample_rate
        self.n_machines = 1 + (len(alt_vad_machine_configs) if alt_vad_machine_configs else 0)
        opt = capi.BatchOptions()
        self._L.fvad_batch_options_default(C.byref(opt))
        opt.device, opt.nn_mode, opt.keep_taps, opt.stft_only = device, nn_mode, int(keep_taps), int(stft_only)
        opt.max_span_chunks = max_span_chunks
        if forced_gains is not None:
            fg = np.ascontiguousarray(forced_gains, np.float32)  # [S*C][frames][161]
            self._keep.append(fg)
            opt.forced_gains = fg.ctypes.data_as(capi.c_f32p)
            opt.forced_gain_frames = fg.shape[1]
        wp = None
        if weights is not None:
            w, k = capi.make_weights(weights)
            self._keep.append(k)
            wp = C.byref(w)
        h = C.c_void_p()
        check(self._L.fvad_batch_create(n_streams, C.byref(self.cfg), wp, C.byref(opt), C.byref(h)))
        self._h = h

    def push(self, pcm: np.ndarray):
        """pcm: host float32 [n_streams*n_channels, n] (stream-major planar)."""
        pcm = np.ascontiguousarray(pcm, np.float32)
        ptrs = (C.c_void_p * pcm.shape[0])(*[pcm[i].ctypes.data for i in range(pcm.shape[0])])
        check(self._L.fvad_batch_push(self._h, ptrs, pcm.shape[0], pcm.shape[1], 0))

    def push_i16(self, pcm: np.ndarray):
        """pcm: host int16 [n_streams*n_channels, n]; converted on the device as x / 32768."""
        pcm = np.ascontiguousarray(pcm, np.int16)
        ptrs = (C.c_void_p * pcm.shape[0])(*[pcm[i].ctypes.data for i in range(pcm.shape[0])])
        check(self._L.fvad_batch_push_i16(self._h, ptrs, pcm.shape[0], pcm.shape[1], 0))

    def push_pointers(self, pointers, n_samples: int, on_device: bool, i16: bool = False):
        """pointers: ints (host or device addresses), one per stream-channel."""
        ptrs = (C.c_void_p * len(pointers))(*pointers)
        fn = self._L.fvad_batch_push_i16 if i16 else self._L.fvad_batch_push
        check(fn(self._h, ptrs, len(pointers), n_samples, 1 if on_device else 0))

    def segments(self, stream: int, machine: int = 0):
        p, n = C.POINTER(capi.Segment)(), C.c_size_t()
        check(self._L.fvad_batch_segments(self._h, stream, machine, C.byref(p), C.byref(n)))
        return _segments_list(p, n.value)

    def tap(self, which: int) -> np.ndarray:
        n = C.c_size_t()
        check(self._L.fvad_batch_tap(self._h, which, None, 0, C.byref(n)))
        dt = {capi.TAP_SPECTRUM: np.complex64, capi.TAP_MET: np.uint8}.get(which, np.float32)
        out = np.empty(n.value, dt)
        check(self._L.fvad_batch_tap(self._h, which, out.ctypes.data, out.nbytes, C.byref(n)))
        sc = self.n_streams * self.n_channels
        if out.size == 0:
            return out.reshape((sc, 0, 161) if which in (capi.TAP_SPECTRUM, capi.TAP_FEATURES, capi.TAP_GAINS) else (sc, 0))
        if which in (capi.TAP_SPECTRUM, capi.TAP_FEATURES, capi.TAP_GAINS):
            return out.reshape(sc, -1, 161)
        if which in (capi.TAP_DENOISED16, capi.TAP_DENOISED):
            return out.reshape(sc, -1)
        return out.reshape(self.n_streams, -1)

    def evaluate(self, references, config: capi.EvalConfig | None = None):
        """references: per stream a list of (from_sec, to_sec).  Returns [stream][machine] dicts of SingleStats."""
        assert len(references) == self.n_streams
        flat = [seg for r in references for seg in r]
        off = (C.c_size_t * (self.n_streams + 1))()
        for s, r in enumerate(references):
            off[s + 1] = off[s] + len(r)
        out = (capi.EvalStats * (self.n_streams * self.n_machines))()
        cfg = config if config is not None else eval_config()
        check(self._L.fvad_batch_evaluate(self._h, _eval_segs(flat), off, C.byref(cfg), out))
        return [[_stats_dict(out[s * self.n_machines + m]) for m in range(self.n_machines)] for s in range(self.n_streams)]

    def stats(self) -> dict:
        st = capi.BatchStats()
        check(self._L.fvad_batch_stats_get(self._h, C.byref(st)))
        return {k: (list(getattr(st, k)) if k.startswith("last_nn_layer") else getattr(st, k)) for k, _ in capi.BatchStats._fields_}

    def close(self):
        if self._h:
            self._L.fvad_batch_destroy(self._h)
            self._h = None

    __del__ = close
