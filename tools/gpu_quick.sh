#!/bin/bash
# tests + bench line only
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err
tail -15 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
