#!/bin/bash
# GPU call 73 (round 2): the driver's round-end sequence on the final tree, the prefill rows, and ncu --set full of the final prefill kernel
bash tools/gpu_call_47.sh
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzl
mkdir -p $O
timeout 600 python tools/prefill_bench.py > $O/prefill.jsonl 2> $O/prefill.err
cut -c1-120 $O/prefill.jsonl
for B in 16 256; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_dequant_tc -s 2 -c 1 -f -o $O/prefill_$B python tools/prefill_one.py $B > $O/ncu_$B.log 2>&1
  tail -1 $O/ncu_$B.log
done
