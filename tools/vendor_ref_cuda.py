#!/usr/bin/env python
"""Vendor + build the reference's OWN CUDA kernels (unmodified) as the "kernel to beat" (SURVEY 8d last row).

    python tools/vendor_ref_cuda.py            # in the build container: /root/reference is present

Copies benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/cuda_kernel.{cu,cpp} byte-for-byte into the git-ignored
baseline/_ref/aqlm_cuda/ (never into the tracked tree) and compiles them there for sm_100a with
torch.utils.cpp_extension (nvcc cross-compiles without a GPU), so the built codebook_cuda.so travels to the GPU box with
the gpurun snapshot.  tools/ref_cuda_bench.py and bench.py import that .so; nothing reads /root/reference at run time.
The reference's own loader (cuda_kernel.py:8-11) uses the same cpp_extension.load call with default flags."""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/benchmark/AQLM/inference_lib/src/aqlm/inference_kernels"
DST = os.path.join(ROOT, "baseline", "_ref", "aqlm_cuda")


def main() -> int:
    if not os.path.isdir(SRC):
        print(f"{SRC} not present (GPU box?): nothing to vendor; using whatever is in {DST}")
        return 0
    os.makedirs(os.path.join(DST, "build"), exist_ok=True)
    for f in ("cuda_kernel.cu", "cuda_kernel.cpp"):
        shutil.copyfile(os.path.join(SRC, f), os.path.join(DST, f))
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", "8")
    from torch.utils.cpp_extension import load
    mod = load(name="codebook_cuda", sources=[os.path.join(DST, "cuda_kernel.cpp"), os.path.join(DST, "cuda_kernel.cu")],
               build_directory=os.path.join(DST, "build"), verbose=True)
    print("built", mod.__file__)
    return 0


if __name__ == "__main__":
    sys.exit(main())
