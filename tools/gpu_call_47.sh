#!/bin/bash
# GPU call 47 (round 2): what the driver runs at round end, on the final tree: pytest -m gpu -x, smoke(), bench, bench --impl reference
cd "$(dirname "$0")/.."
O=gpurun_out/r02zt
mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest.txt 2>&1
tail -4 $O/pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
( time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err ) 2>&1 | grep real
python - <<PY
import json
d=json.loads(open("$O/bench.json").read().strip().splitlines()[-1])
print(d["metric"], d["value"], d["unit"], d["roofline"], d["clocks"], d["gpu_launches"])
print(d["e2e"]["value"], d["decode"]["tok_per_s"], d["cpu_baseline"]["value"])
PY
( time timeout 900 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err ) 2>&1 | grep real
tail -1 $O/bench_ref.json | cut -c1-300
