#!/bin/bash
# GPU call 20 (round 2): full GPU suite with the fused Kx8 path and the pipelined PQ4 kernel; per-format GEMV timing; bench line
cd "$(dirname "$0")/.."
O=gpurun_out/r02t
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 > $O/pytest.txt 2>&1
tail -15 $O/pytest.txt
timeout 600 python tools/pq4_bench.py > $O/pq4_bench.jsonl 2> $O/pq4_bench.err
cat $O/pq4_bench.jsonl | cut -c1-220; tail -3 $O/pq4_bench.err
timeout 600 python tools/kx8_bench.py > $O/kx8_bench.jsonl 2> $O/kx8_bench.err
cat $O/kx8_bench.jsonl; tail -3 $O/kx8_bench.err
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err
tail -2 $O/bench.err
python - <<PY
import json
d=json.loads(open("$O/bench.json").read().strip().splitlines()[-1])
print(d["value"], d["us_per_gemv"], d["dependent_chain_us_per_gemv"], d["roofline"]["frac"])
print(d.get("decode",{}).get("tok_per_s"), d.get("decode_70b",{}).get("tok_per_s"), d["dequant"], d["e2e"]["eager_torch_op_call_us"])
PY
