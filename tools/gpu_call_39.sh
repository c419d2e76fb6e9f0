#!/bin/bash
# GPU call 39 (round 2): ncu capture of a batch-8 grouped GEMV launch (the gather loop in steady state)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zm
mkdir -p $O
timeout 300 python tools/one_batch8.py > $O/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lut_gemv_2x8 -s 6 -c 2 -o $O/batch8 python tools/one_batch8.py > $O/ncu.log 2>&1
tail -3 $O/ncu.log; cat $O/plain.log; ls -la $O
