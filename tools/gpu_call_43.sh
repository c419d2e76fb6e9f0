#!/bin/bash
# GPU call 43 (round 2): ncu capture of the T-MAC qgemm kernel (4096x4096, 4 bits)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zq
mkdir -p $O
timeout 300 python tools/tmac_bench.py > $O/plain.jsonl 2> $O/plain.err && \
ncu --set full --clock-control none --import-source on -k regex:tmac_qgemm_v2 -s 3 -c 1 -o $O/tmac python tools/tmac_bench.py > $O/ncu.log 2>&1
tail -3 $O/ncu.log; ls -la $O
