#!/bin/bash
# GPU call 3 (round 2): set-granular landing slots: parity tests, timing, debug decomposition, timelines with build-phase stamps
cd "$(dirname "$0")/.."
O=gpurun_out/r02c
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -q -x --timeout 120 > $O/pytest.txt 2>&1
tail -5 $O/pytest.txt
N=tools/gemv_bench_dbg
S="--shapes 4096x11008"
{
for T in 8 4; do
  for W in 1 2; do timeout 30 tools/gemv_bench --packed 1 --tuning $T --reps 20 --depworkers $W; done
  timeout 30 tools/gemv_bench --packed 1 --tuning $T --reps 20 --xready
done
for D in 0 0x100 0x200 0x400 0x700; do
  timeout 30 $N --packed 1 --tuning 8 $S --reps 20 --dbg $D --depworkers 1
  timeout 30 $N --packed 1 --tuning 8 $S --reps 20 --dbg $D --depworkers 2
  timeout 30 $N --packed 1 --tuning 8 $S --reps 20 --dbg $D --xready
done
} > $O/times.jsonl 2>&1
cat $O/times.jsonl
timeout 30 $N --timeline --packed 1 --tuning 8 $S --depworkers 1 > $O/tl_dep.txt 2>&1
timeout 30 $N --timeline --packed 1 --tuning 8 $S --xready > $O/tl_xr.txt 2>&1
timeout 30 $N --timeline --packed 1 --tuning 8 --shapes 4096x4096 --depworkers 1 > $O/tl_dep_4096.txt 2>&1
grep -E "L5 " $O/tl_dep.txt $O/tl_xr.txt
