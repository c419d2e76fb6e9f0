#!/bin/bash
# GPU call 5 (round 2): what limits the skeleton?  streaming probe + no-prefetch / no-HBM debug modes
cd "$(dirname "$0")/.."
O=gpurun_out/r02e
mkdir -p $O
timeout 120 tools/stream_probe > $O/stream_probe.jsonl 2>&1
cat $O/stream_probe.jsonl
N=tools/gemv_bench_dbg
S="--shapes 4096x11008"
{
for T in 8,0 8,1; do
for D in 0 0x800 0xF00 0xE00 0xB00 0x700; do
  timeout 30 $N --packed 1 --tuning $T $S --reps 20 --dbg $D --xready
  timeout 30 $N --packed 1 --tuning $T $S --reps 20 --dbg $D --depworkers 2
done
done
} > $O/times.jsonl 2>&1
cat $O/times.jsonl
