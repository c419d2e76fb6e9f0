#!/bin/bash
# GPU call 1 (round 2): microbench check+timing of the rebuilt GEMV, GPU tests, bench line, reference CUDA kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/r02a
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02a/smi.txt
bash tools/gpu_call_a.sh > gpurun_out/r02a/call_a.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/r02a/pytest.txt 2>&1
tail -15 gpurun_out/r02a/pytest.txt
timeout 400 python bench.py > gpurun_out/r02a/bench.json 2> gpurun_out/r02a/bench.err
tail -3 gpurun_out/r02a/bench.err; cat gpurun_out/r02a/bench.json
timeout 400 python tools/ref_cuda_bench.py --out gpurun_out/r02a/kernel_to_beat.json > gpurun_out/r02a/ref_cuda.txt 2>&1
tail -30 gpurun_out/r02a/ref_cuda.txt
cat gpurun_out/r02a/call_a.txt
