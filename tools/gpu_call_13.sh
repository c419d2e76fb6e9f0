#!/bin/bash
# GPU call 13 (round 2): LL exchange on one GPU (multi-rank emulation, decode equivalence), full GPU suite
cd "$(dirname "$0")/.."
O=gpurun_out/r02m
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > $O/pytest.txt 2>&1
tail -15 $O/pytest.txt
