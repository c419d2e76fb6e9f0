#!/bin/bash
# GPU call 8 (round 2): ncu --set full of the GEMV kernel (both streaming modes), one lone launch each
cd "$(dirname "$0")/.."
O=gpurun_out/r02h
mkdir -p $O
B="tools/gemv_bench --ncu --packed 1 --shapes 4096x11008"
$B --tuning 8,1 > $O/plain_bulk.txt 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lut_gemv -s 2 -c 2 -f -o $O/gemv_bulk $B --tuning 8,1 > $O/ncu_bulk.log 2>&1
$B --tuning 8,0 > $O/plain_regs.txt 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lut_gemv -s 2 -c 2 -f -o $O/gemv_regs $B --tuning 8,0 > $O/ncu_regs.log 2>&1
tail -3 $O/ncu_bulk.log $O/ncu_regs.log
ls -la $O
