#!/bin/bash
# GPU call 75 (round 2): the prefill parity tests with four more shapes (single k block, 144 / 80 columns, split-k 16)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzn
mkdir -p $O
timeout -s KILL 400 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "dequant_tensor_core or prefill" > $O/pytest_prefill.txt 2>&1
echo "pytest rc=$?"; tail -25 $O/pytest_prefill.txt
