#!/bin/bash
# GPU call 48 (round 2): bench line of the final tree (cpu_baseline over 16 matrices) + reference arm with the driver's step counts
cd "$(dirname "$0")/.."
O=gpurun_out/r02zu
mkdir -p $O
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err
tail -1 $O/bench.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['frac'], d['cpu_baseline'], d['e2e']['value'])"
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err
tail -1 $O/bench_ref.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('ref', d['value'], d['cpu_baseline']['cores'], d['ms_per_step'])"
