"""The C-ABI library loads on a machine without a GPU, exports every symbol include/fvad.h declares,
its pure helpers work, and every compute entry point fails loudly (FVAD_ERR_NO_DEVICE) instead of
falling back to a CPU path."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


def test_header_symbols_are_exported(fv):
    L = fv.capi.lib()
    hdr = open(os.path.join(ROOT, "include", "fvad.h")).read()
    declared = set(re.findall(r"FVAD_API\s+[\w\s\*]+?\b(fvad_\w+)\s*\(", hdr))
    assert len(declared) >= 30
    assert declared == set(fv.capi.EXPORTS), declared ^ set(fv.capi.EXPORTS)
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/fvad.h but not exported"


def test_pure_helpers_and_defaults(fv):
    L = fv.capi.lib()
    c = fv.vad_machine_config()
    # VADMachine.zig:30-51 defaults
    assert (c.speech_min_freq, c.speech_max_freq, c.long_term_speech_avg_sec) == (500.0, 2000.0, 180.0)
    assert c.has_initial_long_term_avg == 1 and abs(c.initial_long_term_avg - 0.005) < 1e-12
    assert abs(c.short_term_speech_avg_sec - 0.2) < 1e-7 and c.speech_threshold_factor == 10.0
    assert c.channel_vol_ratio_avg_sec == 0.5 and c.channel_vol_ratio_threshold == 0.5
    assert abs(c.min_consecutive_sec_to_open - 0.2) < 1e-7 and c.max_speech_gap_sec == 2.0
    assert abs(c.min_vad_duration_sec - 0.7) < 1e-7
    cfg = fv.capi.Config()
    L.fvad_config_default(C.byref(cfg), 48000, 2)
    assert cfg.sample_rate == 48000 and cfg.n_channels == 2 and cfg.vad_config.fft_size == 1024
    assert L.fvad_nsnet2_chunk_size(48000) == 24000 and L.fvad_nsnet2_chunk_size(16000) == 8000  # NSNet2.zig:157-159
    assert L.fvad_status_name(-1) == b"InvalidSampleRate" and L.fvad_status_name(0) == b"OK"


@pytest.mark.skipif(not _no_gpu(), reason="checks the no-device behaviour")
def test_compute_entry_points_fail_without_device(fv, wd):
    with pytest.raises(fv.FvadError) as e:
        fv.BatchPipeline(1, 48000, 1, wd)
    assert e.value.name == "NoDevice"
    with pytest.raises(fv.FvadError) as e:
        fv.AudioPipeline(48000, 2, weights=wd)
    assert e.value.name == "NoDevice"
    with pytest.raises(fv.FvadError) as e:
        fv.FFT(320, 16000).fft(np.zeros(320, np.float32), np.ones(320, np.float32))
    assert e.value.name == "NoDevice"


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "formula-vad_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dp, f), errors="ignore").read()
                assert "pyoracle" not in src and "fvad_oracle" not in src and "oracle/" not in src, os.path.join(dp, f)
