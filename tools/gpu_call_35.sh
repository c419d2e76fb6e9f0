#!/bin/bash
# GPU call 35 (round 2): grouped rows for the Kx8 formats: parity, timing next to the reference, full GPU suite
cd "$(dirname "$0")/.."
O=gpurun_out/r02zi
mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 300 > $O/pytest.txt 2>&1
tail -6 $O/pytest.txt
timeout 900 python tools/ref_cuda_formats.py > $O/formats.jsonl 2> $O/formats.err
grep "batch 4" $O/formats.jsonl; tail -3 $O/formats.err
