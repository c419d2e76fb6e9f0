#!/bin/bash
# GPU call 34 (round 2): 1x16 GEMV launch-plan sweep
cd "$(dirname "$0")/.."
O=gpurun_out/r02zh
mkdir -p $O
timeout 900 python tools/sweep_1x16.py > $O/sweep.jsonl 2> $O/sweep.err
cat $O/sweep.jsonl; tail -3 $O/sweep.err
