#!/bin/bash
# GPU call 64 (round 2): ncu --set full of the reworked prefill kernel at 16 and 256 rows
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzd
mkdir -p $O
timeout 300 python tools/prefill_one.py 256 > /dev/null 2>&1 || exit 1
for B in 16 256; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_dequant_tc -s 2 -c 1 -f -o $O/prefill_$B python tools/prefill_one.py $B > $O/ncu_$B.log 2>&1
  tail -2 $O/ncu_$B.log
done
ls -la $O
