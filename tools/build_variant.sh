#!/bin/bash
# build_variant.sh NAME [-DFLAG ...] : stand-alone gemv_bench with the library sources compiled in (A/B experiments)
set -e
cd "$(dirname "$0")/.."
NAME=$1; shift
C=rvv-lut-vq_b200/csrc
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ "$@" \
    -o tools/gemv_bench_$NAME tools/gemv_bench.cu $C/api.cu $C/gemv_lut.cu $C/dequant.cu $C/vq_golden.cu $C/gemm_tc.cu $C/shard.cu $C/decode_glue.cu $C/kx8.cu $C/simd_lut.cu $C/tmac.cu $C/code1x16.cu -lcudart_static -ldl -lrt -lpthread
