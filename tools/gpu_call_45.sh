#!/bin/bash
# GPU call 42 (round 2, 4 GPUs): the bench line at N=4 on the final tree (gather parity + sharded 70B decode)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zp
mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 4 --steps 20 --warmup 5 > $O/bench_n4.json 2> $O/bench_n4.err
tail -2 $O/bench_n4.err
python - <<PY
import json
d=json.loads([l for l in open("$O/bench_n4.json").read().strip().splitlines() if l.startswith("{")][-1])
print(d["value"], d["n_gpus"], d.get("parity",{}).get("gather_bit_identical"), d.get("decode_70b",{}).get("tok_per_s"), d.get("decode_70b",{}).get("last_token"))
PY
