#!/bin/bash
# GPU call 31 (round 2, 2 GPUs): multi-GPU parity tests + the bench line at N=2 on the final tree
cd "$(dirname "$0")/.."
O=gpurun_out/r02ze
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_sharded.py -x -m gpu -q --timeout 600 > $O/pytest_2gpu.txt 2>&1
tail -5 $O/pytest_2gpu.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench_n2.json 2> $O/bench_n2.err
tail -2 $O/bench_n2.err
python - <<PY
import json
d=json.loads([l for l in open("$O/bench_n2.json").read().strip().splitlines() if l.startswith("{")][-1])
print(d["value"], d["n_gpus"], d.get("parity"), d.get("decode_70b",{}).get("tok_per_s"), d.get("decode_70b",{}).get("last_token"))
PY
