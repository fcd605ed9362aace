#!/bin/bash
# usage: gpu_ncu.sh <kernel regex> <count> <out name> [bench args]
mkdir -p gpurun_out
RE=$1; CNT=$2; OUT=$3; shift 3
SMALL="python bench.py --hours 0.1 --steps 1 --warmup 1 --no-cpu-baseline --no-48k --e2e-steps 1 $@"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$RE" -c $CNT -o gpurun_out/$OUT $SMALL > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log | cut -c1-300
