#!/bin/bash
# GPU call 26 (round 2): the no-allocation GEMV entry point: its test alone in a fresh process, then the whole GPU suite (it runs last)
cd "$(dirname "$0")/.."
O=gpurun_out/r02z
mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 300 > $O/pytest_noalloc_alone.txt 2>&1
tail -12 $O/pytest_noalloc_alone.txt
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 300 > $O/pytest.txt 2>&1
tail -8 $O/pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
