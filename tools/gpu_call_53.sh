#!/bin/bash
# GPU call 53 (round 2): ncu launch list of the bench step + ncu --set full of the final GEMV kernel
cd "$(dirname "$0")/.."
O=gpurun_out/r02zz
mkdir -p $O
BN="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --no-sweep"
$BN > $O/bench_plain.json 2> $O/bench_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $BN > $O/ncu_launches.log 2>&1
B="tools/gemv_bench --ncu --packed 1 --shapes 4096x11008"
$B --xready > $O/plain_gemv.txt 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lut_gemv -s 2 -c 2 -f -o $O/gemv_final $B --xready > $O/ncu_gemv.log 2>&1
tail -n 3 $O/ncu_gemv.log
ls -la $O
