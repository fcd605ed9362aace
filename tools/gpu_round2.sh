#!/bin/bash
# Round-2 GPU-box visit: parity tests, the bench line, the ncu launch list, a full capture of the top kernels at the small
# size and a full capture of the bench workload's first span (per-launch DRAM traffic for bench.py).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_f16_20x2h.json 2> gpurun_out/r02_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2>> gpurun_out/r02_bench.err
SMALL="python bench.py --hours 0.1 --steps 1 --warmup 1 --no-cpu-baseline --no-48k --e2e-steps 1"
KRE='regex:stft|vadfft|vadscan|tc_|copy_rows|i16_to|upsample|eval_|init_scan|gemm_f32|gru_step'
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$KRE" -c 400 --csv --log-file gpurun_out/r02_launches.csv $SMALL > gpurun_out/ncu_launches.log 2>&1
$SMALL > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:tc_gru|tc_gemm|tc_pack|stft_kernel|istft_|vadfft_kernel|vadscan' -c 11 -o gpurun_out/r02_prof_small $SMALL > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/r02_prof_small.ncu-rep --page raw --csv > gpurun_out/r02_prof_small_raw.csv 2>/dev/null
FULL="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-48k --e2e-steps 1"
ncu --set full --clock-control none -k 'regex:tc_gru3_kernel|tc_gru2_kernel|stft_kernel_d1|istft_stream|tc_gemm_pair_kernel' -c 8 -o gpurun_out/r02_prof_fullsize $FULL > gpurun_out/ncu_fullsize.log 2>&1
ncu -i gpurun_out/r02_prof_fullsize.ncu-rep --page raw --csv > gpurun_out/r02_prof_fullsize_raw.csv 2>/dev/null
tail -3 gpurun_out/r02_pytest_gpu.log; tail -2 gpurun_out/ncu_full.log | cut -c1-200; tail -2 gpurun_out/ncu_fullsize.log | cut -c1-200
python -c "
import json; d=json.load(open('gpurun_out/r02_bench_f16_20x2h.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['cpu_baseline'])"
