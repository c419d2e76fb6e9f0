#!/bin/bash
# GPU call 58 (round 2): prefill kernel with the dequantized A operand in tensor memory (tcgen05.st + A-from-TMEM MMA)
cd "$(dirname "$0")/.."
O=gpurun_out/r02zx
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dequant or prefill or golden_matmat or two_step" > $O/pytest_prefill.txt 2>&1
tail -15 $O/pytest_prefill.txt
LUTVQ_TWO_STEP_MIN_ROWS=100000 timeout 600 python tools/prefill_bench.py > $O/prefill_fused.jsonl 2> $O/prefill_fused.err
cut -c1-130 $O/prefill_fused.jsonl
timeout 600 python tools/prefill_dbg.py 0 1 2 7 39 128 > $O/prefill_dbg.txt 2>&1
tail -20 $O/prefill_dbg.txt
