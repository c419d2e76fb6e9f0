#!/bin/bash
# GPU call 38 (round 2): ncu capture of the 1x16 GEMV (4096->11008)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zl
mkdir -p $O
timeout 300 python tools/one_1x16.py > $O/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:code1x16_gemv -s 6 -c 2 -o $O/code1x16 python tools/one_1x16.py > $O/ncu.log 2>&1
tail -3 $O/ncu.log; cat $O/plain.log; ls -la $O
