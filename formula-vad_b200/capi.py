"""ctypes binding of libfvad_cuda.so (include/fvad.h).  The library has no CPU path: loading
works anywhere, every compute entry point fails with FVAD_ERR_NO_DEVICE without a GPU."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libfvad_cuda.so")

c_f32p = C.POINTER(C.c_float)


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


class VadConfig(C.Structure):
    _fields_ = [("fft_size", C.c_size_t), ("vad_machine_config", VadMachineConfig),
                ("alt_vad_machine_configs", C.POINTER(VadMachineConfig)),
                ("n_alt_vad_machine_configs", C.c_size_t), ("denoiser_model_path", C.c_char_p)]


class Config(C.Structure):
    _fields_ = [("sample_rate", C.c_size_t), ("n_channels", C.c_size_t), ("buffer_length", C.c_size_t),
                ("vad_config", VadConfig), ("skip_processing", C.c_int32)]


WEIGHT_NAMES = ("fc1_w", "fc1_b", "gru1_W", "gru1_R", "gru1_Wb", "gru1_Rb", "gru2_W", "gru2_R",
                "gru2_Wb", "gru2_Rb", "fc2_w", "fc2_b", "fc3_w", "fc3_b", "fc4_w", "fc4_b")


class Weights(C.Structure):
    _fields_ = [("n_bins", C.c_int32), ("hidden", C.c_int32), ("fc2", C.c_int32), ("fc3", C.c_int32),
                ("fc1_relu", C.c_int32)] + [(n, c_f32p) for n in WEIGHT_NAMES]


class BatchOptions(C.Structure):
    _fields_ = [("device", C.c_int32), ("nn_mode", C.c_int32), ("keep_taps", C.c_int32),
                ("stft_only", C.c_int32), ("max_span_chunks", C.c_size_t),
                ("forced_gains", c_f32p), ("forced_gain_frames", C.c_size_t)]


class BatchStats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("chunks_processed", C.c_uint64), ("vad_frames", C.c_uint64),
                ("last_push_device_ms", C.c_double), ("last_stft_ms", C.c_double), ("last_nn_ms", C.c_double),
                ("last_istft_ms", C.c_double), ("last_vadfft_ms", C.c_double), ("last_scan_ms", C.c_double),
                ("last_nn_layer_ms", C.c_double * 9), ("last_nn_layer_launches", C.c_uint32 * 9)]


NN_LAYERS = ("pack", "fc1", "W1", "gru1", "W2", "gru2", "fc2", "fc3", "fc4")


NN_F32, NN_TC_F16X3, NN_TC_F16 = 0, 1, 2
TAP_SPECTRUM, TAP_FEATURES, TAP_GAINS, TAP_DENOISED16, TAP_BAND, TAP_RATIO, TAP_MET, TAP_DENOISED = range(8)

# every symbol include/fvad.h declares (tests check that the library exports each of them)
EXPORTS = (
    "fvad_last_error_string", "fvad_status_name", "fvad_device_count", "fvad_vad_machine_config_default",
    "fvad_config_default", "fvad_fft_create", "fvad_fft_destroy", "fvad_fft_forward", "fvad_fft_inverse",
    "fvad_fft_forward_batch", "fvad_fft_inverse_batch", "fvad_fft_bin_count", "fvad_fft_bin_width",
    "fvad_fft_nyquist_freq", "fvad_fft_freq_to_bin", "fvad_fft_bin_to_freq", "fvad_nsnet2_create",
    "fvad_nsnet2_create_with_weights", "fvad_nsnet2_destroy", "fvad_nsnet2_chunk_size", "fvad_nsnet2_denoise",
    "fvad_nsnet2_calc_spectrogram", "fvad_pipeline_create", "fvad_pipeline_create_with_weights",
    "fvad_pipeline_destroy", "fvad_pipeline_push_samples", "fvad_pipeline_total_write_count",
    "fvad_pipeline_segments", "fvad_batch_options_default", "fvad_batch_create", "fvad_batch_destroy",
    "fvad_batch_push", "fvad_batch_push_i16", "fvad_batch_segments", "fvad_batch_tap", "fvad_batch_stats_get",
    "fvad_eval_config_default", "fvad_evaluate", "fvad_batch_evaluate", "fvad_vad_machine_scan",
    "fvad_vad_machine_scan_noalloc",
)


class Recording(C.Structure):  # fvad_recording (AudioBuffer handed to AudioPipeline.Callbacks)
    _fields_ = [("n_channels", C.c_size_t), ("length", C.c_size_t), ("sample_rate", C.c_size_t),
                ("global_start_frame_number", C.c_uint64), ("channel_pcm", C.POINTER(C.POINTER(C.c_float)))]


RECORDING_CB = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(Recording))


class Callbacks(C.Structure):  # fvad_callbacks
    _fields_ = [("ctx", C.c_void_p), ("on_original_recording", RECORDING_CB), ("on_denoised_recording", RECORDING_CB)]


class EvalSegment(C.Structure):  # fvad_eval_segment
    _fields_ = [("from_sec", C.c_float), ("to_sec", C.c_float)]


class EvalConfig(C.Structure):  # fvad_eval_config
    _fields_ = [("ignore_shorter_than_sec", C.c_float), ("extrude_start", C.c_float), ("extrude_end", C.c_float),
                ("fill_gaps", C.c_float), ("f_score_beta", C.c_float)]


class EvalStats(C.Structure):  # fvad_eval_stats
    _fields_ = [(n, C.c_float) for n in ("total_positives_sec", "true_positives_sec", "false_positives_sec",
                                          "false_negatives_sec", "true_positive_rate", "precision",
                                          "false_negative_rate", "false_discovery_rate", "f_score", "fm_index")]


class FvadError(RuntimeError):
    def __init__(self, code: int, name: str, msg: str):
        super().__init__(f"{name} ({code}): {msg}")
        self.code, self.name = code, name


_lib = None


def lib():
    """Load the CUDA library.  Raises if it has not been built: there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} is missing: run ./build.sh (no CPU fallback exists)")
    L = C.CDLL(LIB_PATH)
    vp, sz, i32 = C.c_void_p, C.c_size_t, C.c_int

    def sig(name, res, *args):
        f = getattr(L, name)
        f.restype, f.argtypes = res, list(args)

    sig("fvad_last_error_string", C.c_char_p)
    sig("fvad_status_name", C.c_char_p, i32)
    sig("fvad_device_count", i32, C.POINTER(i32))
    sig("fvad_vad_machine_config_default", None, C.POINTER(VadMachineConfig))
    sig("fvad_config_default", None, C.POINTER(Config), sz, sz)
    sig("fvad_fft_create", i32, sz, sz, i32, C.POINTER(vp))
    sig("fvad_fft_destroy", None, vp)
    sig("fvad_fft_forward", i32, vp, c_f32p, sz, c_f32p, sz, c_f32p, sz, C.POINTER(Complex), sz)
    sig("fvad_fft_inverse", i32, vp, C.POINTER(Complex), sz, c_f32p, sz)
    sig("fvad_fft_forward_batch", i32, vp, vp, sz, sz, vp, vp, i32)
    sig("fvad_fft_inverse_batch", i32, vp, vp, sz, vp, i32)
    sig("fvad_fft_bin_count", sz, vp)
    sig("fvad_fft_bin_width", C.c_float, vp)
    sig("fvad_fft_nyquist_freq", C.c_float, vp)
    sig("fvad_fft_freq_to_bin", i32, vp, C.c_float, C.POINTER(sz))
    sig("fvad_fft_bin_to_freq", i32, vp, sz, c_f32p)
    sig("fvad_nsnet2_create", i32, sz, C.c_char_p, C.POINTER(vp))
    sig("fvad_nsnet2_create_with_weights", i32, sz, C.POINTER(Weights), i32, C.POINTER(vp))
    sig("fvad_nsnet2_destroy", None, vp)
    sig("fvad_nsnet2_chunk_size", sz, sz)
    sig("fvad_nsnet2_denoise", i32, vp, c_f32p, sz, c_f32p, sz, c_f32p, sz)
    sig("fvad_nsnet2_calc_spectrogram", i32, c_f32p, sz, sz, c_f32p, sz, C.POINTER(Complex), sz)
    sig("fvad_pipeline_create", i32, C.POINTER(Config), vp, C.POINTER(vp))
    sig("fvad_pipeline_create_with_weights", i32, C.POINTER(Config), vp, C.POINTER(Weights), i32, C.POINTER(vp))
    sig("fvad_pipeline_destroy", None, vp)
    sig("fvad_pipeline_push_samples", i32, vp, C.POINTER(c_f32p), sz, sz, C.POINTER(C.c_uint64))
    sig("fvad_pipeline_total_write_count", C.c_uint64, vp)
    sig("fvad_pipeline_segments", i32, vp, sz, C.POINTER(C.POINTER(Segment)), C.POINTER(sz))
    sig("fvad_batch_options_default", None, C.POINTER(BatchOptions))
    sig("fvad_batch_create", i32, sz, C.POINTER(Config), C.POINTER(Weights), C.POINTER(BatchOptions), C.POINTER(vp))
    sig("fvad_batch_destroy", None, vp)
    sig("fvad_batch_push", i32, vp, C.POINTER(vp), sz, sz, i32)
    sig("fvad_batch_push_i16", i32, vp, C.POINTER(vp), sz, sz, i32)
    sig("fvad_batch_segments", i32, vp, sz, sz, C.POINTER(C.POINTER(Segment)), C.POINTER(sz))
    sig("fvad_batch_tap", i32, vp, i32, vp, sz, C.POINTER(sz))
    sig("fvad_batch_stats_get", i32, vp, C.POINTER(BatchStats))
    sig("fvad_eval_config_default", None, C.POINTER(EvalConfig))
    sig("fvad_evaluate", i32, C.POINTER(EvalSegment), sz, C.POINTER(EvalSegment), sz, C.POINTER(EvalConfig), C.POINTER(EvalStats))
    sig("fvad_batch_evaluate", i32, vp, C.POINTER(EvalSegment), C.POINTER(sz), C.POINTER(EvalConfig), C.POINTER(EvalStats))
    sig("fvad_vad_machine_scan", i32, C.POINTER(VadMachineConfig), sz, sz, sz, c_f32p, c_f32p, sz, sz, vp, C.POINTER(Segment), sz,
        C.POINTER(sz))
    sig("fvad_vad_machine_scan_noalloc", i32, C.POINTER(VadMachineConfig), sz, sz, sz, vp, vp, sz, vp, vp, sz, vp, vp)
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        L = lib()
        raise FvadError(rc, L.fvad_status_name(rc).decode(), L.fvad_last_error_string().decode())


def make_weights(wd: dict):
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
