#!/bin/bash
# GPU call 21 (round 2): PQ4 byte-table kernel: parity (both kernels forced) + throughput
cd "$(dirname "$0")/.."
O=gpurun_out/r02u
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_pq4.py -m gpu -q --timeout 300 > $O/pytest_pq4.txt 2>&1
tail -15 $O/pytest_pq4.txt
timeout 600 python tools/pq4_bench.py > $O/pq4_bench.jsonl 2> $O/pq4_bench.err
cat $O/pq4_bench.jsonl | cut -c1-330; tail -3 $O/pq4_bench.err
