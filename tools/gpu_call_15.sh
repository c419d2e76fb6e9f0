#!/bin/bash
# GPU call 15 (round 2, --gpus 8): bench.py at N=8 (parity + weak-scaling GEMV ring + 70B LL decode), 70B decode at N=4
cd "$(dirname "$0")/.."
O=gpurun_out/r02o
mkdir -p $O
nvidia-smi -L > $O/gpus.txt
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 900 $TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench_n8.json 2> $O/bench_n8.err
tail -3 $O/bench_n8.err
python - <<PY
import json
try:
    txt=open("$O/bench_n8.json").read(); d=json.loads([l for l in txt.split("\n") if l.startswith("{")][-1])
    print(d["value"], d["us_per_gemv"], d.get("parity"), d.get("decode_70b"))
except Exception as e: print("no bench line", e)
PY
CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 600 $TR --nproc-per-node 4 --master-port 29522 tools/decode_bench.py --model 70b --exchange ll > $O/decode70_n4_ll.json 2> $O/decode70_n4.err
tail -2 $O/decode70_n4.err; cat $O/decode70_n4_ll.json
timeout 600 $TR --nproc-per-node 8 --master-port 29523 tools/decode_bench.py --model 7b --exchange ll > $O/decode7_n8_ll.json 2> $O/decode7_n8.err
cat $O/decode7_n8_ll.json
