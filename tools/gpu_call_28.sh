#!/bin/bash
# GPU call 28 (round 2): rows of a multi-row GEMV call as one grouped launch: parity tests, small-batch timing
cd "$(dirname "$0")/.."
O=gpurun_out/r02zb
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -x -m gpu -q --timeout 300 > $O/pytest_rows.txt 2>&1
tail -12 $O/pytest_rows.txt
timeout 600 python tools/small_batch_bench.py > $O/small_batch.jsonl 2> $O/small_batch.err
cat $O/small_batch.jsonl; tail -3 $O/small_batch.err
