#!/bin/bash
# GPU call 46 (round 2): T-MAC qgemm with half-block items + the 1x16 plan-override test
cd "$(dirname "$0")/.."
O=gpurun_out/r02zs
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_tmac.py tests/test_gpu_1x16.py -m gpu -q --timeout 300 > $O/pytest.txt 2>&1
tail -5 $O/pytest.txt
timeout 600 python tools/tmac_bench.py > $O/tmac.jsonl 2> $O/tmac.err
cat $O/tmac.jsonl; tail -3 $O/tmac.err
