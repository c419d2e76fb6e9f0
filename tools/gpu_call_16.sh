#!/bin/bash
# GPU call 16 (round 2): full GPU suite (HF loader, dequant with 4 loads in flight), bench line
cd "$(dirname "$0")/.."
O=gpurun_out/r02p
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > $O/pytest.txt 2>&1
tail -6 $O/pytest.txt
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err
tail -2 $O/bench.err
python - <<PY
import json
d=json.load(open("$O/bench.json"))
print(d["value"], d["us_per_gemv"], d["dependent_chain_us_per_gemv"], d["roofline"]["frac"])
print(d.get("decode",{}).get("tok_per_s"), d.get("decode_70b",{}).get("tok_per_s"), d["dequant"], d["e2e"]["eager_torch_op_call_us"])
PY
