#!/bin/bash
# GPU call 23 (round 2): full GPU suite (prefill dispatch, PQ4 wide-LDS kernel) + bench line
cd "$(dirname "$0")/.."
O=gpurun_out/r02w
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q --timeout 300 > $O/pytest.txt 2>&1
tail -15 $O/pytest.txt
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err
tail -2 $O/bench.err
python - <<PY
import json
d=json.loads(open("$O/bench.json").read().strip().splitlines()[-1])
print(d["value"], d["us_per_gemv"], d["dependent_chain_us_per_gemv"], d["roofline"]["frac"])
print(d.get("decode",{}).get("tok_per_s"), d.get("decode_70b",{}).get("tok_per_s"), d["dequant"], d["e2e"]["eager_torch_op_call_us"])
for r in d["prefill"]: print(r)
PY
