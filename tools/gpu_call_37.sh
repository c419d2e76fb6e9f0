#!/bin/bash
# GPU call 37 (round 2): dependent GEMV launches with 3 workers per SM (80-register variant) vs 2 (128-register variant): 7B decode
cd "$(dirname "$0")/.."
O=gpurun_out/r02zk
mkdir -p $O
run() { LUTVQ_DEP_WORKERS=$1 timeout 300 python tools/decode_bench.py --model 7b --plain $2 2>> $O/decode.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('workers=$1 $2', round(d['value'],1), round(d['us_per_token'],1))"; }
run 2 ""
run 3 "--tuning 8,1,1"
run 2 "--tuning 8,1,1"
run 3 ""
run 1 "--tuning 8,1,1"
tail -3 $O/decode.err
