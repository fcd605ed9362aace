"""formula-vad_b200: B200-native drop-in for Formula-VAD's per-frame audio hot path.

The directory name carries a hyphen (the task names it so); import it through the repo-root
helper `fvad_b200.py`, which registers this package as `formula_vad_b200`.
The compute lives in csrc/ (hand-written sm_100a CUDA behind the C ABI of include/fvad.h);
this package is the thin host-side mirror of the reference's AudioPipeline / FFT / NSNet2 surface.
"""
from . import capi, sharding, synth, weights  # noqa: F401
from .pipeline import (FFT, NSNet2, AudioPipeline, BatchPipeline, Evaluator, VADMachine, FvadError, eval_config,  # noqa: F401
                       vad_machine_config)
