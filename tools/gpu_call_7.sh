#!/bin/bash
# GPU call 7 (round 2): pair-word atomics + global scale restored on top of the hot-line fixes (x slice via smem, one coalesced
# norm pass, codebook read once); trigger-at-entry experiment
cd "$(dirname "$0")/.."
O=gpurun_out/r02g
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > $O/pytest.txt 2>&1
tail -8 $O/pytest.txt
N=tools/gemv_bench_dbg
S="--shapes 4096x11008"
{
for T in 8,0 8,1; do
  for W in 1 2; do timeout 30 tools/gemv_bench --packed 1 --tuning $T --reps 20 --depworkers $W; done
  timeout 30 tools/gemv_bench --packed 1 --tuning $T --reps 20 --xready
done
for T in 8,0 8,1; do
for D in 0 0x100 0x400 0x700 0x1000 0x1700 0x2000; do
  timeout 30 $N --packed 1 --tuning $T $S --reps 20 --dbg $D --xready
  timeout 30 $N --packed 1 --tuning $T $S --reps 20 --dbg $D --depworkers 2
done
done
} > $O/times.jsonl 2>&1
cat $O/times.jsonl
timeout 30 $N --timeline --packed 1 --tuning 8,1 $S --depworkers 1 > $O/tl_dep_bulk.txt 2>&1
timeout 30 $N --timeline --packed 1 --tuning 8,0 $S --xready > $O/tl_xr_regs.txt 2>&1
timeout 30 $N --timeline --packed 1 --tuning 8,1 $S --xready > $O/tl_xr_bulk.txt 2>&1
grep -E "L5 " $O/tl_dep_bulk.txt $O/tl_xr_regs.txt $O/tl_xr_bulk.txt
