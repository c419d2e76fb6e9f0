#!/bin/bash
# GPU call 40 (round 2): pending-atomic ring 4 deep (variant 8,2,4) vs 2 deep: small batches and the 7B decoder
cd "$(dirname "$0")/.."
O=gpurun_out/r02zn
mkdir -p $O
for t in "8,2,2" "8,2,4"; do
  echo "tuning $t"
  GEMV_TUNING=$t timeout 300 python tools/small_batch_bench.py 2>> $O/err.log | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(' ', d['shape'], d['batch'], d['grouped_us'])"
  timeout 300 python tools/decode_bench.py --model 7b --plain --tuning $t 2>> $O/err.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('  decode', round(d['value'],1), round(d['us_per_token'],1))"
done
tail -3 $O/err.log
