#!/bin/bash
# GPU call 33 (round 2): 1x16 GEMV with L1-bypassing codebook reads and 16-row CTAs: parity + timing next to the reference's kernel
cd "$(dirname "$0")/.."
O=gpurun_out/r02zg
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_1x16.py -m gpu -q --timeout 300 > $O/pytest_1x16.txt 2>&1
tail -4 $O/pytest_1x16.txt
timeout 900 python tools/ref_cuda_formats.py > $O/formats.jsonl 2> $O/formats.err
cat $O/formats.jsonl; tail -3 $O/formats.err
