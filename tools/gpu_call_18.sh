#!/bin/bash
# GPU call 18 (round 2): restored kernel, independent ring with / without the L2 prefetch of the share, ring depths
cd "$(dirname "$0")/.."
O=gpurun_out/r02r
mkdir -p $O
N=tools/gemv_bench_dbg
{
for T in 8,1,1 8,2,2 8,2,8; do
for D in 0 0x800; do
  timeout 30 $N --packed 1 --tuning $T --reps 20 --dbg $D --xready
done
done
timeout 30 $N --packed 1 --reps 20 --dbg 0x800
timeout 30 $N --packed 1 --reps 20
} > $O/times.jsonl 2>&1
cat $O/times.jsonl
