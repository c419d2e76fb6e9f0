#!/bin/bash
# GPU call 27 (round 2): rows of one multi-row GEMV call overlapped: parity tests, small-batch timing, full suite, bench line
cd "$(dirname "$0")/.."
O=gpurun_out/r02za
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 300 -k "multi_row or matmat or linear" > $O/pytest_rows.txt 2>&1
tail -8 $O/pytest_rows.txt
timeout 600 python tools/small_batch_bench.py > $O/small_batch.jsonl 2> $O/small_batch.err
cat $O/small_batch.jsonl; tail -3 $O/small_batch.err
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 300 > $O/pytest.txt 2>&1
tail -6 $O/pytest.txt
timeout 900 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err
tail -2 $O/bench.err
python - <<PY
import json
d=json.loads(open("$O/bench.json").read().strip().splitlines()[-1])
print(d["value"], d["us_per_gemv"], d["dependent_chain_us_per_gemv"], d["roofline"]["frac"])
print(d.get("decode",{}).get("tok_per_s"), d.get("decode_70b",{}).get("tok_per_s"), d["dequant"], d["e2e"]["eager_torch_op_call_us"], d["e2e"].get("eager_torch_op_only_us"))
PY
