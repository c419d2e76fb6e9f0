#!/bin/bash
# GPU call 82 (round 2): split-k final reduction with 16-byte loads of 4 rows x 4 columns per thread: parity, then the shapes against HEAD in the same call
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzu
mkdir -p $O
C=rvv-lut-vq_b200/csrc
timeout -s KILL 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dequant or prefill or golden_matmat or two_step" > $O/pytest_prefill.txt 2>&1
echo "pytest rc=$?"; tail -3 $O/pytest_prefill.txt
cp $C/liblutvq.so $C/liblutvq_new.so
echo new; timeout -s KILL 300 python tools/prefill_shapes.py 2>&1 | cut -c1-110
cp $C/liblutvq_base.so $C/liblutvq.so
echo base; timeout -s KILL 300 python tools/prefill_shapes.py 2>&1 | head -6 | cut -c1-110
cp $C/liblutvq_new.so $C/liblutvq.so
