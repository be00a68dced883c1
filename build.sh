#!/bin/bash
# Builds zuds_pipeline_b200/libzb200.so (hand-written sm_100a CUDA + C ABI) in-tree.
# nvcc cross-compiles without a GPU.  -fmad=false on the files whose discrete decisions and fit
# arithmetic must match the reference's unfused x86 doubles bit for bit; the convolution pass
# keeps FMA contraction and uses explicit __dmul_rn/__dadd_rn where order matters.
# Any failed compile fails the build: objects are removed first and every job's status is checked.
set -e
cd "$(dirname "$0")/zuds_pipeline_b200/csrc"
ARCH="-gencode arch=compute_100a,code=sm_100a"
COMMON="-O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-ffp-contract=off $ARCH"
mkdir -p ../../build
NOFMAD="zb_host zb_stamps zb_fit zb_stats zb_tile zb_bkg zb_align zb_check"
FMAD="zb_mask zb_conv zb_prepost"
pids=()
for f in $NOFMAD $FMAD; do rm -f ../../build/$f.o; done
for f in $NOFMAD; do nvcc $COMMON -fmad=false -c $f.cu -o ../../build/$f.o & pids+=($!); done
for f in $FMAD;   do nvcc $COMMON             -c $f.cu -o ../../build/$f.o & pids+=($!); done
fail=0
for p in "${pids[@]}"; do wait $p || fail=1; done
if [ $fail -ne 0 ]; then echo "build.sh: compilation failed" >&2; exit 1; fi
OBJS=""
for f in $NOFMAD $FMAD; do OBJS="$OBJS ../../build/$f.o"; done
nvcc $ARCH -shared -o ../libzb200.so $OBJS -lcudart
echo "built zuds_pipeline_b200/libzb200.so"
