#!/bin/bash
# GPU call 71 (round 2, 2 GPUs): GPU tests with two devices visible (the sharded-decode tests run) and the bench line at N=2, final tree
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzj
mkdir -p $O
timeout 1200 python -m pytest tests -x -q -m gpu > $O/pytest_2gpu.txt 2>&1
tail -3 $O/pytest_2gpu.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err
tail -2 $O/bench_n2.err
python - <<PY
import json
d=json.loads([l for l in open("$O/bench_n2.json").read().strip().splitlines() if l.startswith("{")][-1])
print(d["value"], d["n_gpus"], d.get("parity",{}).get("gather_bit_identical"), d.get("decode_70b",{}).get("tok_per_s"), d.get("decode_70b",{}).get("last_token"))
PY
