#!/bin/bash
# GPU call 30 (round 2, 2 GPUs): sharded 70B decode, LL exchange with and without dataflow GEMV launches
cd "$(dirname "$0")/.."
O=gpurun_out/r02zd
mkdir -p $O
for v in "--no-dataflow" ""; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/decode_bench.py --model 70b --exchange ll $v 2>> $O/decode.err | grep '^{' | tee -a $O/decode.jsonl | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['exchange'], d['dataflow'], round(d['value'],1), round(d['us_per_token'],1), d['launches_per_token'], d['wait_errors'])"
done
tail -3 $O/decode.err
