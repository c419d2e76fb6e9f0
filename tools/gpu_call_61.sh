#!/bin/bash
# GPU call 61 (round 2): prefill kernel: chunk-0 code loads before the codebook, programmatic dependent launch (mode 2048 = off)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zza
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dequant or prefill or golden_matmat or two_step" > $O/pytest_prefill.txt 2>&1
tail -5 $O/pytest_prefill.txt
timeout 600 python tools/prefill_dbg.py 0 2048 39 > $O/prefill_dbg.txt 2>&1
tail -20 $O/prefill_dbg.txt
