"""ctypes binding of the CPU ORACLE (oracle/_build/libfvad_oracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Nothing in the product package imports this file.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_build", "libfvad_oracle.so")
REF_KISS_PATH = os.path.join(HERE, "_ref", "libkissfft_ref.so")

c_f32p = C.POINTER(C.c_float)
c_f64p = C.POINTER(C.c_double)


class Complex(C.Structure):
    _fields_ = [("r", C.c_float), ("i", C.c_float)]


class Segment(C.Structure):
    _fields_ = [("sample_from", C.c_uint64), ("sample_to", C.c_uint64),
                ("avg_channel_vol_ratio", C.c_float), ("vad_met_sec", C.c_float)]


class VadMachineConfig(C.Structure):
    _fields_ = [("speech_min_freq", C.c_float), ("speech_max_freq", C.c_float),
                ("long_term_speech_avg_sec", C.c_float), ("has_initial_long_term_avg", C.c_int32),
                ("initial_long_term_avg", C.c_double), ("short_term_speech_avg_sec", C.c_float),
                ("speech_threshold_factor", C.c_float), ("channel_vol_ratio_avg_sec", C.c_float),
                ("channel_vol_ratio_threshold", C.c_float), ("min_consecutive_sec_to_open", C.c_float),
                ("max_speech_gap_sec", C.c_float), ("min_vad_duration_sec", C.c_float)]

    @classmethod
    def default(cls, **kw):
        """Defaults of src/AudioPipeline/VADMachine.zig:30-51."""
        c = cls(500, 2000, 180, 1, 0.005, 0.2, 10, 0.5, 0.5, 0.2, 2, 0.7)
        for k, v in kw.items():
            setattr(c, k, v)
        return c


class VadConfig(C.Structure):
    _fields_ = [("fft_size", C.c_size_t), ("vad_machine_config", VadMachineConfig),
                ("alt_vad_machine_configs", C.POINTER(VadMachineConfig)),
                ("n_alt_vad_machine_configs", C.c_size_t), ("denoiser_model_path", C.c_char_p)]


class Config(C.Structure):
    _fields_ = [("sample_rate", C.c_size_t), ("n_channels", C.c_size_t), ("buffer_length", C.c_size_t),
                ("vad_config", VadConfig), ("skip_processing", C.c_int32)]


class Weights(C.Structure):
    _fields_ = [("n_bins", C.c_int32), ("hidden", C.c_int32), ("fc2", C.c_int32), ("fc3", C.c_int32),
                ("fc1_relu", C.c_int32)] + [(n, c_f32p) for n in (
                    "fc1_w", "fc1_b", "gru1_W", "gru1_R", "gru1_Wb", "gru1_Rb",
                    "gru2_W", "gru2_R", "gru2_Wb", "gru2_Rb",
                    "fc2_w", "fc2_b", "fc3_w", "fc3_b", "fc4_w", "fc4_b")]


WEIGHT_NAMES = ("fc1_w", "fc1_b", "gru1_W", "gru1_R", "gru1_Wb", "gru1_Rb", "gru2_W", "gru2_R",
                "gru2_Wb", "gru2_Rb", "fc2_w", "fc2_b", "fc3_w", "fc3_b", "fc4_w", "fc4_b")


def make_weights_struct(wd: dict):
    """wd: dict name -> np.float32 array (see formula-vad_b200 weights module).  Returns
    (struct, keepalive)."""
    w = Weights()
    w.n_bins, w.hidden = wd["fc1_w"].shape[1], wd["fc1_w"].shape[0]
    w.fc2, w.fc3 = wd["fc2_w"].shape[0], wd["fc3_w"].shape[0]
    w.fc1_relu = int(wd.get("fc1_relu", 0))
    keep = []
    for n in WEIGHT_NAMES:
        a = np.ascontiguousarray(wd[n], dtype=np.float32)
        keep.append(a)
        setattr(w, n, a.ctypes.data_as(c_f32p))
    return w, keep


class EvalSeg(C.Structure):
    _fields_ = [("from_sec", C.c_float), ("to_sec", C.c_float)]


class EvalCfg(C.Structure):
    _fields_ = [("extrude_start", C.c_float), ("extrude_end", C.c_float), ("fill_gaps", C.c_float),
                ("ignore_shorter_than", C.c_float), ("beta", C.c_float)]


class EvalStats(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("p_sec", "tp_sec", "fp_sec", "fn_sec", "tpr", "ppv", "fnr",
                                          "fdr", "f_score", "fm")]


class Meta(C.Structure):  # AudioPipeline/VADMetadata.zig accumulator
    _fields_ = [("has_ratio", C.c_int), ("ratio_sum", C.c_float), ("ratio_weight", C.c_float),
                ("has_min", C.c_int), ("vmin", C.c_float), ("has_max", C.c_int), ("vmax", C.c_float)]


class MetaResult(C.Structure):
    _fields_ = [("has_ratio", C.c_int), ("ratio", C.c_float), ("has_min", C.c_int), ("vmin", C.c_float),
                ("has_max", C.c_int), ("vmax", C.c_float)]


def build(force: bool = False) -> None:
    """Compile the oracle (and oracle/_ref when the reference checkout is present)."""
    if force or not os.path.exists(LIB_PATH) or any(
            os.path.getmtime(os.path.join(HERE, f)) > os.path.getmtime(LIB_PATH)
            for f in os.listdir(HERE) if f.endswith((".c", ".h"))):
        subprocess.check_call(["make", "-C", HERE, "_build/libfvad_oracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/lib/kissfft") and (force or not os.path.exists(REF_KISS_PATH)):
        subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p

    def sig(name, res, *args):
        f = getattr(L, name)
        f.restype = res
        f.argtypes = list(args)

    sig("orc_rfft_new", vp, C.c_int, C.c_int)
    sig("orc_rfft_free", None, vp)
    sig("orc_rfft_forward", None, vp, c_f32p, C.POINTER(Complex))
    sig("orc_rfft_inverse", None, vp, C.POINTER(Complex), c_f32p)
    sig("orc_use_kissfft", None, vp, vp, vp)
    sig("orc_hann_periodic", None, c_f32p, C.c_int)
    sig("orc_hann_symmetric", None, c_f32p, C.c_int)
    sig("orc_window_norm_factor", C.c_float, c_f32p, C.c_int)
    sig("orc_nsnet2_window", None, c_f32p)
    sig("orc_rms_volume", C.c_float, c_f32p, C.c_size_t)
    sig("orc_upsample", C.c_float, c_f32p, C.c_size_t, c_f32p, C.c_float, C.c_int)
    sig("orc_freq_to_bin", C.c_int, C.c_float, C.c_size_t, C.c_size_t)
    sig("orc_fftop_new", vp, C.c_int, C.c_int)
    sig("orc_fftop_free", None, vp)
    sig("orc_fftop_forward", None, vp, c_f32p, c_f32p, C.POINTER(Complex))
    sig("orc_fftop_inverse", None, vp, C.POINTER(Complex), c_f32p)
    sig("orc_nn_new", vp, C.POINTER(Weights))
    sig("orc_nn_free", None, vp)
    sig("orc_nn_forward", None, vp, c_f32p, c_f32p, C.c_int)
    sig("orc_nn_forward_f64", None, C.POINTER(Weights), c_f32p, c_f64p, C.c_int)
    sig("orc_nsnet2_new", vp, C.c_size_t, C.POINTER(Weights))
    sig("orc_nsnet2_free", None, vp)
    sig("orc_nsnet2_chunk_size", C.c_size_t, C.c_size_t)
    sig("orc_nsnet2_denoise", C.c_int, vp, c_f32p, C.c_size_t, c_f32p)
    sig("orc_nsnet2_last_spec", C.POINTER(Complex), vp)
    sig("orc_nsnet2_last_features", c_f32p, vp)
    sig("orc_nsnet2_last_gains", c_f32p, vp)
    sig("orc_nsnet2_force_gains", None, vp, c_f32p)
    sig("orc_rolling_new", vp, C.c_size_t, C.c_int, C.c_double)
    sig("orc_rolling_free", None, vp)
    sig("orc_rolling_push", C.c_double, vp, C.c_float)
    sig("orc_meta_push", None, C.POINTER(Meta), C.POINTER(MetaResult), C.c_float)
    sig("orc_meta_to_result", MetaResult, C.POINTER(Meta))
    sig("orc_downsample", None, c_f32p, C.c_size_t, c_f32p, C.c_int)
    sig("orc_vadm_new", vp, C.POINTER(VadMachineConfig), C.c_size_t, C.c_size_t, C.c_size_t)
    sig("orc_vadm_free", None, vp)
    sig("orc_vadm_run", C.c_int, vp, c_f32p, C.c_uint64, C.c_int, C.c_float, C.POINTER(C.c_uint64))
    sig("orc_vadm_segments", C.c_size_t, vp, C.POINTER(C.POINTER(Segment)))
    sig("orc_vadm_last_met", C.c_int, vp)
    sig("orc_vadm_state", C.c_int, vp)
    sig("orc_pipeline_new", vp, C.POINTER(Config), C.POINTER(Weights), C.c_int, C.POINTER(C.c_int))
    sig("orc_pipeline_free", None, vp)
    sig("orc_pipeline_push", C.c_int, vp, C.POINTER(c_f32p), C.c_size_t, C.POINTER(C.c_uint64))
    sig("orc_pipeline_total_write_count", C.c_uint64, vp)
    sig("orc_pipeline_segments", C.c_size_t, vp, C.c_size_t, C.POINTER(C.POINTER(Segment)))
    sig("orc_pipeline_frames", C.c_size_t, vp, C.POINTER(c_f32p), C.POINTER(c_f32p), C.POINTER(C.POINTER(C.c_uint8)))
    sig("orc_pipeline_keep_denoised", None, vp, C.c_int)
    sig("orc_pipeline_denoised", C.c_size_t, vp, C.c_size_t, C.POINTER(c_f32p))
    sig("orc_pipeline_keep_nn_taps", None, vp, C.c_int)
    sig("orc_pipeline_nn_taps", C.c_size_t, vp, C.POINTER(c_f32p), C.POINTER(c_f32p), C.POINTER(C.POINTER(Complex)))
    sig("orc_pipeline_force_gains", None, vp, c_f32p, C.c_size_t)
    sig("orc_evaluate", None, C.POINTER(EvalSeg), C.c_size_t, C.POINTER(EvalSeg), C.c_size_t,
        C.POINTER(EvalCfg), C.POINTER(EvalStats))
    sig("orc_eval_false_positive_sec", C.c_float, C.POINTER(EvalSeg), C.POINTER(EvalSeg), C.c_size_t,
        C.POINTER(EvalCfg))
    _lib = L
    return L


def f32p(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_f32p)


_kiss = None


def use_reference_kissfft(on: bool = True) -> bool:
    """Route the oracle's real FFTs through oracle/_ref/libkissfft_ref.so (the reference's own
    kissfft C, compiled unmodified).  Returns False if that library is not present."""
    global _kiss
    L = lib()
    if not on:
        L.orc_use_kissfft(None, None, None)
        return True
    if not os.path.exists(REF_KISS_PATH):
        return False
    if _kiss is None:
        _kiss = C.CDLL(REF_KISS_PATH)
    g = lambda n: C.cast(getattr(_kiss, n), C.c_void_p)
    L.orc_use_kissfft(g("kiss_fftr_alloc"), g("kiss_fftr"), g("kiss_fftri"))
    return True


def rfft(x: np.ndarray, window: np.ndarray | None = None) -> np.ndarray:
    """Windowed forward real FFT exactly as src/FFT.zig:85-113 (batch over leading dims)."""
    L = lib()
    x = np.ascontiguousarray(x, np.float32)
    n = x.shape[-1]
    w = np.ones(n, np.float32) if window is None else np.ascontiguousarray(window, np.float32)
    flat = x.reshape(-1, n)
    out = np.empty((flat.shape[0], n // 2 + 1), np.complex64)
    op = L.orc_fftop_new(n, 0)
    for i in range(flat.shape[0]):
        L.orc_fftop_forward(op, f32p(flat[i]), f32p(w), out[i].ctypes.data_as(C.POINTER(Complex)))
    L.orc_fftop_free(op)
    return out.reshape(x.shape[:-1] + (n // 2 + 1,))


def irfft(X: np.ndarray, n: int) -> np.ndarray:
    """Unnormalised inverse real FFT as src/FFT.zig:115-134 (kiss_fftri)."""
    L = lib()
    X = np.ascontiguousarray(X, np.complex64)
    flat = X.reshape(-1, n // 2 + 1)
    out = np.empty((flat.shape[0], n), np.float32)
    op = L.orc_fftop_new(n, 1)
    for i in range(flat.shape[0]):
        L.orc_fftop_inverse(op, flat[i].ctypes.data_as(C.POINTER(Complex)), f32p(out[i]))
    L.orc_fftop_free(op)
    return out.reshape(X.shape[:-1] + (n,))


def nsnet2_window() -> np.ndarray:
    w = np.empty(320, np.float32)
    lib().orc_nsnet2_window(f32p(w))
    return w


def hann_periodic(n: int) -> np.ndarray:
    w = np.empty(n, np.float32)
    lib().orc_hann_periodic(f32p(w), n)
    return w


def nn_forward(wd: dict, features: np.ndarray, f64: bool = False) -> np.ndarray:
    """features [steps,161] -> gains [steps,161], zero initial state."""
    L = lib()
    ws, keep = make_weights_struct(wd)
    f = np.ascontiguousarray(features, np.float32)
    steps = f.shape[0]
    if f64:
        g = np.empty((steps, f.shape[1]), np.float64)
        L.orc_nn_forward_f64(C.byref(ws), f32p(f), g.ctypes.data_as(c_f64p), steps)
        return g
    g = np.empty((steps, f.shape[1]), np.float32)
    nn = L.orc_nn_new(C.byref(ws))
    L.orc_nn_forward(nn, f32p(f), f32p(g), steps)
    L.orc_nn_free(nn)
    return g


def make_config(sample_rate=48000, n_channels=1, main: VadMachineConfig | None = None, alts=None,
                fft_size=1024):
    cfg = Config()
    cfg.sample_rate, cfg.n_channels, cfg.buffer_length, cfg.skip_processing = sample_rate, n_channels, 0, 0
    cfg.vad_config.fft_size = fft_size
    cfg.vad_config.vad_machine_config = main or VadMachineConfig.default()
    keep = None
    if alts:
        keep = (VadMachineConfig * len(alts))(*alts)
        cfg.vad_config.alt_vad_machine_configs = keep
        cfg.vad_config.n_alt_vad_machine_configs = len(alts)
    return cfg, keep


class Pipeline:
    """AudioPipeline-shaped wrapper over the oracle (src/AudioPipeline.zig:40-143)."""

    def __init__(self, wd: dict | None, sample_rate=48000, n_channels=1, main=None, alts=None,
                 allow_any_rate=True, keep_denoised=False, keep_nn=False, forced_gains=None):
        L = lib()
        self.L = L
        self.cfg, self._k1 = make_config(sample_rate, n_channels, main, alts)
        self.n_channels = n_channels
        self.n_machines = 1 + (len(alts) if alts else 0)
        if wd is not None:
            self.ws, self._k2 = make_weights_struct(wd)
            wp = C.byref(self.ws)
        else:
            wp = None
        err = C.c_int(0)
        self.h = L.orc_pipeline_new(C.byref(self.cfg), wp, int(allow_any_rate), C.byref(err))
        if not self.h:
            raise ValueError({-1: "InvalidSampleRate", -2: "InvalidFFTSize"}.get(err.value, str(err.value)))
        if keep_denoised:
            L.orc_pipeline_keep_denoised(self.h, 1)
        if keep_nn:
            L.orc_pipeline_keep_nn_taps(self.h, 1)
        if forced_gains is not None:
            self._fg = np.ascontiguousarray(forced_gains, np.float32)  # [C][F][161]
            L.orc_pipeline_force_gains(self.h, f32p(self._fg), self._fg.shape[1])

    def push_samples(self, pcm: np.ndarray) -> int:
        """pcm: [n_channels, n] float32 planar."""
        pcm = np.ascontiguousarray(pcm, np.float32)
        assert pcm.shape[0] == self.n_channels
        ptrs = (c_f32p * self.n_channels)(*[f32p(pcm[c]) for c in range(self.n_channels)])
        first = C.c_uint64(0)
        rc = self.L.orc_pipeline_push(self.h, ptrs, pcm.shape[1], C.byref(first))
        if rc:
            raise RuntimeError(f"oracle push failed: {rc}")
        return first.value

    def segments(self, machine=0):
        p = C.POINTER(Segment)()
        n = self.L.orc_pipeline_segments(self.h, machine, C.byref(p))
        return [(p[i].sample_from, p[i].sample_to, p[i].avg_channel_vol_ratio, p[i].vad_met_sec) for i in range(n)]

    def frames(self):
        b, r, m = c_f32p(), c_f32p(), C.POINTER(C.c_uint8)()
        n = self.L.orc_pipeline_frames(self.h, C.byref(b), C.byref(r), C.byref(m))
        if n == 0:
            return np.zeros(0, np.float32), np.zeros(0, np.float32), np.zeros(0, np.uint8)
        return (np.ctypeslib.as_array(b, (n,)).copy(), np.ctypeslib.as_array(r, (n,)).copy(),
                np.ctypeslib.as_array(m, (n,)).copy())

    def denoised(self, ch=0):
        p = c_f32p()
        n = self.L.orc_pipeline_denoised(self.h, ch, C.byref(p))
        return np.ctypeslib.as_array(p, (n,)).copy() if n else np.zeros(0, np.float32)

    def nn_taps(self):
        f, g, s = c_f32p(), c_f32p(), C.POINTER(Complex)()
        n = self.L.orc_pipeline_nn_taps(self.h, C.byref(f), C.byref(g), C.byref(s))
        if n == 0:
            return None
        feat = np.ctypeslib.as_array(f, (n, 161)).copy()
        gain = np.ctypeslib.as_array(g, (n, 161)).copy()
        spec = np.ctypeslib.as_array(C.cast(s, c_f32p), (n, 161, 2)).copy().view(np.complex64)[..., 0]
        return feat, gain, spec

    def close(self):
        if self.h:
            self.L.orc_pipeline_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
