#!/bin/bash
# GPU call A (round 2): correctness + timing sweep of the rebuilt LUT-GEMV through the C-ABI microbenchmark
cd "$(dirname "$0")/.."
O=gpurun_out/r02a
mkdir -p $O
B=tools/gemv_bench
for T in 16,2 8,4 8,2; do
  timeout 120 $B --tuning $T --check --reps 5 --packed 1 > $O/check_${T/,/_}.txt 2>&1
  timeout 120 $B --tuning $T --check --reps 5 --packed 1 --xready >> $O/check_${T/,/_}.txt 2>&1
done
timeout 120 $B --check --reps 5 --packed 0 > $O/check_ref_layout.txt 2>&1
for T in 16,2 8,4 8,2; do
  timeout 120 $B --tuning $T --reps 30 --packed 1 --xready > $O/time_xready_${T/,/_}.jsonl 2>&1
  timeout 120 $B --tuning $T --reps 30 --packed 1 > $O/time_dep_${T/,/_}.jsonl 2>&1
done
timeout 120 $B --reps 20 --packed 0 > $O/time_ref_layout.jsonl 2>&1
tail -n +1 $O/check_*.txt | grep -v "^{" | head -60
cat $O/time_*.jsonl
