#!/bin/bash
# GPU call 55 (round 2): prefill kernel with the activation tile brought in by a TMA tensor copy: parity, then the prefill rows
cd "$(dirname "$0")/.."
O=gpurun_out/r02zu
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dequant or prefill or golden_matmat or two_step" > $O/pytest_prefill.txt 2>&1
tail -5 $O/pytest_prefill.txt
timeout 600 python tools/prefill_bench.py > $O/prefill.jsonl 2> $O/prefill.err
cut -c1-260 $O/prefill.jsonl
tail -3 $O/prefill.err
