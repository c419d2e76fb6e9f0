#!/bin/bash
# GPU call 2 (round 2): full GPU tests after the staging fix; timelines + debug decomposition of the new GEMV kernel and of
# the round-1 kernel (built from commit 6cf8acd) on the same box; new bench.py end to end
cd "$(dirname "$0")/.."
O=gpurun_out/r02b
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > $O/pytest.txt 2>&1
tail -15 $O/pytest.txt
N=tools/gemv_bench_dbg
R=tools/gemv_bench_r1dbg
S="--shapes 4096x11008"
for T in 8,2 16,2; do
  timeout 60 $N --timeline --packed 1 --tuning $T $S > $O/tl_new_dep_${T/,/_}.txt 2>&1
  timeout 60 $N --timeline --packed 1 --tuning $T $S --xready > $O/tl_new_xr_${T/,/_}.txt 2>&1
done
timeout 60 $N --timeline --packed 1 --tuning 8,2 $S --depworkers 2 > $O/tl_new_dep2w_8_2.txt 2>&1
timeout 60 $R --timeline --packed 1 $S > $O/tl_r1_dep.txt 2>&1
timeout 60 $R --timeline --packed 1 $S --xready > $O/tl_r1_xr.txt 2>&1
{
for D in 0 0x100 0x200 0x400 0x300 0x700; do
  timeout 60 $N --packed 1 --tuning 8,2 $S --reps 20 --dbg $D
  timeout 60 $N --packed 1 --tuning 8,2 $S --reps 20 --dbg $D --xready
  timeout 60 $N --packed 1 --tuning 16,2 $S --reps 20 --dbg $D --xready
done
for W in 1 2; do
  timeout 60 tools/gemv_bench --packed 1 --tuning 8,2 --reps 20 --depworkers $W
  timeout 60 tools/gemv_bench --packed 1 --tuning 16,2 --reps 20 --depworkers $W
  timeout 60 tools/gemv_bench --packed 1 --tuning 8,4 --reps 20 --depworkers $W
done
timeout 60 $R --packed 1 --reps 20
timeout 60 $R --packed 1 --reps 20 --xready
for D in 0x100 0x200 0x400 0x700; do
  timeout 60 $R --packed 1 $S --reps 20 --dbg $D
  timeout 60 $R --packed 1 $S --reps 20 --dbg $D --xready
done
} > $O/dbg_times.jsonl 2>&1
cat $O/dbg_times.jsonl
timeout 600 python bench_new.py --steps 20 --warmup 5 > $O/bench_new.json 2> $O/bench_new.err
tail -5 $O/bench_new.err
head -c 3000 $O/bench_new.json
