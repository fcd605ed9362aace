#!/bin/bash
mkdir -p gpurun_out
FVAD_STFT_STREAM=1 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_stft_stream.log 2>&1; echo "pytest(stream) rc=$?" >> gpurun_out/pytest_gpu_stft_stream.log
tail -3 gpurun_out/pytest_gpu_stft_stream.log
bash tools/gpu_ab.sh FVAD_STFT_STREAM=1 FVAD_AB=tile FVAD_STFT_STREAM=1 FVAD_AB=tile2
SMALL="python bench.py --nn-mode f16 --hours 0.1 --steps 1 --warmup 1 --no-cpu-baseline --no-48k --e2e-steps 1"
FVAD_STFT_STREAM=1 ncu --set full --clock-control none --import-source on -k 'regex:stft_' -c 2 -o gpurun_out/prof_stft_stream $SMALL > gpurun_out/ncu_stft_stream.log 2>&1
tail -2 gpurun_out/ncu_stft_stream.log | cut -c1-200
