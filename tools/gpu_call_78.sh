#!/bin/bash
# GPU call 78 (round 2): chained prefill launches (each call's input = the previous call's output) against the same chain with synchronisations
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzq
mkdir -p $O
timeout -s KILL 400 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "prefill_chain" > $O/pytest_chain.txt 2>&1
echo "pytest rc=$?"; tail -25 $O/pytest_chain.txt
