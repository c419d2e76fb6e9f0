#!/bin/bash
# GPU call 41 (round 2): full GPU suite + bench line on the tree with the 4-deep pending ring for dependent launches
cd "$(dirname "$0")/.."
O=gpurun_out/r02zo
mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 300 > $O/pytest.txt 2>&1
tail -5 $O/pytest.txt
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err
tail -2 $O/bench.err
python - <<PY
import json
d=json.loads(open("$O/bench.json").read().strip().splitlines()[-1])
print(d["value"], d["us_per_gemv"], d["dependent_chain_us_per_gemv"], d["roofline"]["frac"])
print(d.get("decode",{}).get("tok_per_s"), d.get("decode_70b",{}).get("tok_per_s"), d["dequant"], d["e2e"]["eager_torch_op_call_us"], d["e2e"].get("eager_torch_op_only_us"))
print(d.get("small_batch"))
PY
