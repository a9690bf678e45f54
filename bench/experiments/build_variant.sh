#!/bin/bash
# build_variant.sh NAME [-DFLAG ...]: the native library compiled with extra preprocessor flags into
# build/variants/NAME/libpoisson_b200.so (select it at run time with POISSON_B200_LIB=...).  For A/B runs of kernel
# variants inside one gpurun call: boxes of the pool differ by up to 30 % in latency-bound numbers, so only
# numbers taken in the same call on the same box compare.
set -e
NAME=$1; shift
ROOT=$(cd "$(dirname "$0")/../.." && pwd)
OUT=$ROOT/build/variants/$NAME
mkdir -p "$OUT"
SRCS="core/decomp.cpp cpu/oracle.cpp cpu/dist_solver.cpp runtime/shm_comm.cpp api/c_api_cpu.cpp cuda/assemble.cu cuda/step_kernels.cu cuda/persistent.cu cuda/persistent_tma.cu cuda/resident.cu cuda/apply_a_op.cu runtime/solver.cu runtime/solver_run.cu runtime/group.cu api/c_api_cuda.cu"
pids=()
for s in $SRCS; do
  o=$OUT/$(echo $s | tr / _).o
  nvcc -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -rdc=false -gencode arch=compute_100a,code=sm_100a \
    -Xcompiler -O3,-fPIC,-fopenmp,-ffp-contract=off -I "$ROOT/include" -I "$ROOT/csrc" "$@" -c "$ROOT/csrc/$s" -o "$o" &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o "$OUT/libpoisson_b200.so" $OUT/*.o -Xcompiler -fopenmp -lgomp
rm -f $OUT/*.o
echo "$OUT/libpoisson_b200.so"
