#!/bin/bash
# GPU call 81 (round 2): per-CTA timeline of the prefill kernel on the split-k shape 4096x4096 (instrumented build)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzt
mkdir -p $O
timeout -s KILL 300 python tools/prefill_timeline.py 4096 4096 128 > $O/timeline_4096.txt 2>&1
cat $O/timeline_4096.txt | tail -12
