"""Kernel selection with the reference's function names (aqlm/inference_kernels/kernel_selector.py:21-145).

The reference dispatches over CUDA / Triton / numba / generic-dequantize back ends.  Here there is exactly one back end:
the sm_100a kernels of liblutvq behind `torch.ops.aqlm.*`.  Formats with a kernel return it; anything else raises
NotImplementedError -- no Triton, no CPU route, no silent generic fallback (deliberate: DESIGN.md "no compatibility layers")."""
from __future__ import annotations

import warnings
from contextlib import contextmanager
from typing import Callable

import torch

from . import torch_ops  # noqa: F401  (registers torch.ops.aqlm.*)


@contextmanager
def optimize_for_training():
    """kernel_selector.py:8-18: deprecated no-op kept for source compatibility."""
    warnings.warn("`optimize_for_training` is deprecated. The optimization now happens automatically at runtime.")
    yield


def _format(codebooks: torch.Tensor):
    num_codebooks, codebook_size, out_group_size, in_group_size = codebooks.shape
    if codebooks.device.type != "cuda":
        raise NotImplementedError("lutvq has no CPU / XPU path: move the module to a CUDA device (sm_100a)")
    if out_group_size != 1:
        raise NotImplementedError(f"out_group_size {out_group_size} has no lutvq kernel (the reference routes it to Triton / generic dequantize)")
    if (num_codebooks, codebook_size) == (1, 65536) and in_group_size in (8, 16):
        return "code1x16"
    if codebook_size == 256 and in_group_size == 8:
        return {2: "code2x8", 1: "code1x8"}.get(num_codebooks, "codeKx8" if num_codebooks <= 8 else None)
    return None


def get_forward_pass_kernel(codebooks: torch.Tensor, optimize_for_training: bool) -> Callable:
    """kernel_selector.py:21-102: False -> the GEMV-class op (`*_matmat`), True -> the dequantize + GEMM op (`*_matmat_dequant`)."""
    fmt = _format(codebooks)
    if fmt is None:
        raise NotImplementedError(f"no lutvq kernel for codebooks of shape {tuple(codebooks.shape)}")
    if fmt == "codeKx8":
        return torch.ops.aqlm.codeKx8_matmat                 # N = 3..8 x 256: one op for both batch regimes
    return getattr(torch.ops.aqlm, fmt + ("_matmat_dequant" if optimize_for_training else "_matmat"))


def get_backward_pass_kernel(codebooks: torch.Tensor, optimize_for_training: bool) -> Callable:
    """kernel_selector.py:105-145: grad_input = (grad_output * scales) @ W through the `*_matmat_dequant_transposed` ops."""
    fmt = _format(codebooks)
    if fmt in ("code1x16", "code2x8", "code1x8"):
        return getattr(torch.ops.aqlm, fmt + "_matmat_dequant_transposed")
    shape = tuple(codebooks.shape)

    def _no_backward(*args, **kwargs):               # selected lazily like the reference's ops: only a backward CALL fails
        raise NotImplementedError(f"no lutvq backward kernel for codebooks of shape {shape}")
    return _no_backward
