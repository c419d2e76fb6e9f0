#!/bin/bash
# GPU call 57 (round 2): prefill kernel decomposition after the TMA change
cd "$(dirname "$0")/.."
O=gpurun_out/r02zw
mkdir -p $O
timeout 600 python tools/prefill_dbg.py 0 1 2 4 7 39 55 128 > $O/prefill_dbg.txt 2>&1
cat $O/prefill_dbg.txt | tail -30
