#!/bin/bash
# One GPU-box visit after a spectral-kernel change: parity tests, A/B of the ISTFT variants, ncu capture of the spectral kernels.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
bash tools/gpu_ab.sh FVAD_ISTFT_CHUNK=1 FVAD_AB=stream FVAD_ISTFT_CHUNK=1 FVAD_AB=stream2
SMALL="python bench.py --nn-mode f16 --hours 0.1 --steps 1 --warmup 1 --no-cpu-baseline --no-48k --e2e-steps 1"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:stft_kernel|istft_' -c 4 -o gpurun_out/prof_spec $SMALL > gpurun_out/ncu_spec.log 2>&1
tail -2 gpurun_out/ncu_spec.log | cut -c1-200
