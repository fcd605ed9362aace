#!/bin/bash
# Builds formula-vad_b200/libfvad_cuda.so (sm_100a only) in-tree.  Usage: ./build.sh [-j N]
set -e
cd "$(dirname "$0")/formula-vad_b200/csrc"
OUT=../libfvad_cuda.so
OBJ=../_obj
mkdir -p $OBJ
NVCC=${NVCC:-nvcc}
FLAGS="-std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -Xcompiler -fPIC,-fvisibility=hidden -I../../include"
pids=()
for f in *.cu *.cpp; do
  o=$OBJ/${f%.*}.o
  if [ ! -f $o ] || [ $f -nt $o ] || [ -n "$(find . ../../include -name '*.cuh' -newer $o -o -name '*.h' -newer $o | head -1)" ]; then
    ( $NVCC $FLAGS -c $f -o $o ) &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -o $OUT $OBJ/*.o -lcudart -lcuda
echo "built $(realpath $OUT)"
