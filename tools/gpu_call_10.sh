#!/bin/bash
# GPU call 10 (round 2): 32-bit shared loads + IDP.2A instead of 16-bit shared loads in the gather
cd "$(dirname "$0")/.."
O=gpurun_out/r02j
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -q -x --timeout 120 > $O/pytest.txt 2>&1
tail -4 $O/pytest.txt
N=tools/gemv_bench_dbg
S="--shapes 4096x11008"
{
for T in 8,0 8,1; do
  for W in 1 2; do timeout 30 tools/gemv_bench --packed 1 --tuning $T --reps 20 --depworkers $W; done
  timeout 30 tools/gemv_bench --packed 1 --tuning $T --reps 20 --xready
done
for D in 0x100 0x400 0x700; do
  timeout 30 $N --packed 1 --tuning 8,1 $S --reps 20 --dbg $D --xready
  timeout 30 $N --packed 1 --tuning 8,1 $S --reps 20 --dbg $D --depworkers 2
done
} > $O/times.jsonl 2>&1
cat $O/times.jsonl
timeout 30 $N --timeline --packed 1 --tuning 8,1 $S --depworkers 1 > $O/tl_dep_bulk.txt 2>&1
timeout 30 $N --timeline --packed 1 --tuning 8,1 $S --xready > $O/tl_xr_bulk.txt 2>&1
grep -E "L5 " $O/tl_dep_bulk.txt $O/tl_xr_bulk.txt
