#!/bin/bash
# GPU call 56 (round 2): prefill kernel with a TMA warp of its own (copies run up to a whole ring ahead): parity, prefill rows by
# dispatch and with the fused kernel forced at every batch
cd "$(dirname "$0")/.."
O=gpurun_out/r02zv
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dequant or prefill or golden_matmat or two_step" > $O/pytest_prefill.txt 2>&1
tail -5 $O/pytest_prefill.txt
timeout 600 python tools/prefill_bench.py > $O/prefill.jsonl 2> $O/prefill.err
cut -c1-130 $O/prefill.jsonl
LUTVQ_TWO_STEP_MIN_ROWS=100000 timeout 600 python tools/prefill_bench.py > $O/prefill_fused.jsonl 2> $O/prefill_fused.err
cut -c1-130 $O/prefill_fused.jsonl
tail -3 $O/prefill.err $O/prefill_fused.err
