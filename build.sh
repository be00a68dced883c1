#!/bin/bash
# Builds zuds_pipeline_b200/libzb200.so (hand-written sm_100a CUDA + C ABI) in-tree.
# nvcc cross-compiles without a GPU.  -fmad=false on the files whose discrete decisions and fit
# arithmetic must match the reference's unfused x86 doubles bit for bit; the convolution pass
# keeps FMA contraction and uses explicit __dmul_rn/__dadd_rn where order matters.
set -e
cd "$(dirname "$0")/zuds_pipeline_b200/csrc"
ARCH="-gencode arch=compute_100a,code=sm_100a"
COMMON="-O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-ffp-contract=off $ARCH"
mkdir -p ../../build
nvcc $COMMON -fmad=false -c zb_host.cu   -o ../../build/zb_host.o &
nvcc $COMMON             -c zb_mask.cu   -o ../../build/zb_mask.o &
nvcc $COMMON -fmad=false -c zb_stamps.cu -o ../../build/zb_stamps.o &
nvcc $COMMON -fmad=false -c zb_fit.cu    -o ../../build/zb_fit.o &
nvcc $COMMON             -c zb_conv.cu   -o ../../build/zb_conv.o &
nvcc $COMMON -fmad=false -c zb_stats.cu  -o ../../build/zb_stats.o &
nvcc $COMMON -fmad=false -c zb_tile.cu   -o ../../build/zb_tile.o &
nvcc $COMMON -fmad=false -c zb_bkg.cu    -o ../../build/zb_bkg.o &
wait
nvcc $ARCH -shared -o ../libzb200.so ../../build/zb_host.o ../../build/zb_mask.o ../../build/zb_stamps.o \
     ../../build/zb_fit.o ../../build/zb_conv.o ../../build/zb_stats.o ../../build/zb_tile.o ../../build/zb_bkg.o -lcudart
echo "built zuds_pipeline_b200/libzb200.so"
