#!/bin/bash
# GPU call 36 (round 2): T-MAC bit-serial GEMV: parity tests + timing
cd "$(dirname "$0")/.."
O=gpurun_out/r02zj
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_tmac.py -m gpu -q --timeout 300 > $O/pytest_tmac.txt 2>&1
tail -5 $O/pytest_tmac.txt
timeout 600 python tools/tmac_bench.py > $O/tmac.jsonl 2> $O/tmac.err
cat $O/tmac.jsonl; tail -3 $O/tmac.err
