#!/bin/bash
# GPU call 12 (round 2): x slice through shared memory (one request per hot line and CTA) on the restored kernel
cd "$(dirname "$0")/.."
O=gpurun_out/r02l
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -q -x --timeout 120 > $O/pytest.txt 2>&1
tail -4 $O/pytest.txt
{
timeout 30 tools/gemv_bench --packed 1 --reps 20
timeout 30 tools/gemv_bench --packed 1 --reps 20 --xready
} > $O/times.jsonl 2>&1
cat $O/times.jsonl
timeout 600 python bench.py --steps 20 --warmup 5 --no-extras --no-cpu-baseline > $O/bench.json 2> $O/bench.err
tail -2 $O/bench.err
python - <<PY
import json
d=json.load(open("$O/bench.json"))
print(d["value"], d["us_per_gemv"], d["dependent_chain_us_per_gemv"], d["roofline"]["frac"], [ (r["shape"], r["x_ready"]["us"], r["dependent"]["us"]) for r in d["sweep"]])
PY
