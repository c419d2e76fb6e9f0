#!/bin/bash
# GPU call 84 (round 2): prefill kernel, the 4 MMAs of a k block behind one election, against HEAD in the same call
cd "$(dirname "$0")/.."
O=gpurun_out/r02zzw
mkdir -p $O
C=rvv-lut-vq_b200/csrc
timeout -s KILL 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dequant or prefill or golden_matmat or two_step" > $O/pytest_prefill.txt 2>&1
echo "pytest rc=$?"; tail -3 $O/pytest_prefill.txt
cp $C/liblutvq.so $C/liblutvq_new.so
for rep in 1 2; do
echo new; timeout -s KILL 300 python tools/prefill_dbg.py 0 2>&1 | tail -4
cp $C/liblutvq_base.so $C/liblutvq.so
echo base; timeout -s KILL 300 python tools/prefill_dbg.py 0 2>&1 | tail -4
cp $C/liblutvq_new.so $C/liblutvq.so
done
