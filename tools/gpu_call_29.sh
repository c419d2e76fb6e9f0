#!/bin/bash
# GPU call 29 (round 2): LL decode with dataflow GEMV launches on ONE GPU: parity, then tok/s of the variants (7B)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zc
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_sharded.py -x -m gpu -q --timeout 300 > $O/pytest_sharded.txt 2>&1
tail -12 $O/pytest_sharded.txt
for v in "--plain" "--no-dataflow" ""; do
  timeout 300 python tools/decode_bench.py --model 7b --exchange ll $v 2>> $O/decode.err | tee -a $O/decode.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['exchange'], d['dataflow'], round(d['value'],1), round(d['us_per_token'],1), d['launches_per_token'], d['wait_errors'])"
done
tail -3 $O/decode.err
