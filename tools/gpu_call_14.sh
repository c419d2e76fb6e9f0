#!/bin/bash
# GPU call 14 (round 2, --gpus 2): multi-GPU parity test log, bench.py at N=2, 70B decode ll vs push
cd "$(dirname "$0")/.."
O=gpurun_out/r02n
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -q --timeout 300 -k "multi_gpu" > $O/pytest_2gpu.txt 2>&1
tail -5 $O/pytest_2gpu.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err
tail -3 $O/bench_n2.err
python - <<PY
import json
try:
    d=json.load(open("$O/bench_n2.json"))
    print(d["value"], d["us_per_gemv"], d.get("parity"), d.get("decode_70b"))
except Exception as e: print("no bench line", e)
PY
for X in ll push; do
timeout 600 $TR --master-port 29512 tools/decode_bench.py --model 70b --exchange $X > $O/decode70_$X.json 2> $O/decode70_$X.err
tail -2 $O/decode70_$X.err; cat $O/decode70_$X.json
done
