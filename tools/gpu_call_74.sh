#!/bin/bash
# GPU call 74 (round 2): final prefill kernel, k blocks per stage revisited (mode 256 = 2, 128 = 1; default 4 / 2 / 1 by batch)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzm
mkdir -p $O
timeout -s KILL 300 python tools/prefill_dbg.py 0 256 128 0 > $O/bps.txt 2>&1
cat $O/bps.txt | tail -16
