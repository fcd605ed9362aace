#!/bin/bash
# same-box A/B of two builds of the library: gpu_ab_lib.sh build/libA.so build/libB.so [repeat]
L=formula-vad_b200/libfvad_cuda.so
cp $L /tmp/keep.so
for r in $(seq 1 ${3:-2}); do
for lib in $1 $2; do
  cp $lib $L
  python bench.py --steps 2 --warmup 2 --hours 1 --no-cpu-baseline --no-48k --e2e-steps 1 > gpurun_out/ab2.json 2>gpurun_out/ab2.err
  python - <<PY
import json
d=json.load(open("gpurun_out/ab2.json"))
print("$lib", "value=%.0f"%d["value"], {k:round(v,2) for k,v in d["stage_ms_per_step"].items()}, "pack", round(d["nn_layer_ms_per_step"]["pack"],2), "e2e16 %.0f"%d["e2e_i16"]["value"])
PY
done; done
cp /tmp/keep.so $L
