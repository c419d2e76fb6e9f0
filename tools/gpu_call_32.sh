#!/bin/bash
# GPU call 32 (round 2): the reference's CUDA kernels for 1x16 / 1x8 next to ours
cd "$(dirname "$0")/.."
O=gpurun_out/r02zf
mkdir -p $O
timeout 900 python tools/ref_cuda_formats.py > $O/formats.jsonl 2> $O/formats.err
cat $O/formats.jsonl; tail -3 $O/formats.err
