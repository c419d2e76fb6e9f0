#!/bin/bash
# GPU call 51 (round 2): more shapes through the grouped small-batch path (ragged rows, few k-slices, tiny layers)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/r02zx
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 300 -k "multi_row" > gpurun_out/r02zx/pytest.txt 2>&1
tail -15 gpurun_out/r02zx/pytest.txt
