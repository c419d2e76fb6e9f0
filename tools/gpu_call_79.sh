#!/bin/bash
# GPU call 79 (round 2): prefill routes on the other 7B shapes
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzr
mkdir -p $O
timeout -s KILL 600 python tools/prefill_shapes.py > $O/shapes.jsonl 2> $O/shapes.err
cat $O/shapes.jsonl; tail -3 $O/shapes.err
