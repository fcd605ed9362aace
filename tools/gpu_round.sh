#!/bin/bash
# One GPU-box visit: parity tests, the bench line, ncu launch list, a full capture of the top kernels at the small
# size (source pages) and a full capture of the bench workload's first span (per-launch DRAM traffic for bench.py).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py --nn-mode f16 --steps 3 --warmup 3 > gpurun_out/bench_f16.json 2> gpurun_out/bench_f16.err
SMALL="python bench.py --nn-mode f16 --hours 0.1 --steps 1 --warmup 1 --no-cpu-baseline --no-48k --e2e-steps 1"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:stft|vadfft|vadscan|tc_|copy_rows|i16_to|upsample|eval_|init_scan|gemm_f32|gru_step' -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
$SMALL > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:tc_gru|tc_gemm|stft_kernel|istft_|vadfft_kernel|vadscan' -c 14 -o gpurun_out/prof_r01 $SMALL > gpurun_out/ncu_full.log 2>&1
FULL="python bench.py --nn-mode f16 --steps 1 --warmup 1 --no-cpu-baseline --no-48k --e2e-steps 1"
ncu --set full --clock-control none -k 'regex:tc_gru_kernel|stft_kernel_d1|istft_stream' -c 4 -o gpurun_out/prof_r01_fullsize $FULL > gpurun_out/ncu_fullsize.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_f16.json
