#!/bin/bash
# GPU call 70 (round 2): the driver's round-end sequence on the tree with the converged MMA issue + the prefill rows by tools/prefill_dbg.py
bash tools/gpu_call_47.sh
cd "$(dirname "$0")/.."
timeout -s KILL 300 python tools/prefill_dbg.py 0 2>&1 | tail -4
