#!/bin/bash
# GPU call 49 (round 2): write-only bandwidth probe
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/r02zv
timeout 300 python tools/write_bw.py 2>&1 | tee gpurun_out/r02zv/write_bw.jsonl
