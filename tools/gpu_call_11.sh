#!/bin/bash
# GPU call 11 (round 2): round-1 kernel restored (fp32-pattern LUT of 64 tables, per-slice 2^e scales, register rings) + L2 hints
cd "$(dirname "$0")/.."
O=gpurun_out/r02k
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > $O/pytest.txt 2>&1
tail -8 $O/pytest.txt
{
timeout 30 tools/gemv_bench --packed 1 --reps 20
timeout 30 tools/gemv_bench --packed 1 --reps 20 --xready
timeout 30 tools/gemv_bench --packed 0 --reps 20
} > $O/times.jsonl 2>&1
cat $O/times.jsonl
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err
tail -2 $O/bench.err
python - <<PY
import json
d=json.load(open("$O/bench.json"))
print(d["value"], d["us_per_gemv"], d["dependent_chain_us_per_gemv"], d["roofline"]["frac"], [ (r["shape"], r["x_ready"]["us"], r["dependent"]["us"]) for r in d["sweep"]])
print(d.get("decode",{}).get("tok_per_s"), d.get("decode_70b",{}).get("tok_per_s"), d["e2e"])
PY
