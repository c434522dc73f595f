#!/bin/bash
# Builds libb2t_sm100.so in-tree for sm_100a (cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --use_fast_math -Xptxas -v"
# --use_fast_math would relax erff/division accuracy: keep IEEE paths, fast intrinsics are opted in per call
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC"
mkdir -p build
OBJS=""
pids=""
for f in *.cu; do
  o=build/${f%.cu}.o
  OBJS="$OBJS $o"
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ] || [ common.cuh -nt "$o" ] || [ ../../include/b2t_sm100.h -nt "$o" ]; then
    ( $NVCC $FLAGS ${PTXAS_V:+-Xptxas -v} -c "$f" -o "$o" 2> build/${f%.cu}.log || { cat build/${f%.cu}.log; exit 1; } ) &
    pids="$pids $!"
  fi
done
# second pass over gemm_tc.cu: the 1-term instantiations (see B2T_GEMM_PART in the file)
o=build/gemm_tc_part1.o
OBJS="$OBJS $o"
if [ ! -f "$o" ] || [ gemm_tc.cu -nt "$o" ] || [ common.cuh -nt "$o" ] || [ ../../include/b2t_sm100.h -nt "$o" ]; then
  ( $NVCC $FLAGS -DB2T_GEMM_PART=1 ${PTXAS_V:+-Xptxas -v} -c gemm_tc.cu -o "$o" 2> build/gemm_tc_part1.log || { cat build/gemm_tc_part1.log; exit 1; } ) &
  pids="$pids $!"
fi
for p in $pids; do wait $p || exit 1; done
$NVCC -shared -o libb2t_sm100.so $OBJS -cudart static
echo "built $(pwd)/libb2t_sm100.so"
