#!/bin/bash
# GPU call 68 (round 2): per-CTA timeline of the prefill kernel
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzh
mkdir -p $O
timeout -s KILL 300 python tools/prefill_timeline.py > $O/timeline.txt 2>&1
cat $O/timeline.txt | tail -12
