"""`torch.ops.aqlm.*` registration: the reference's operator names and schema
(benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/cuda_kernel.py:13-132: torch.library.define / impl / impl_abstract for
code1x16 / code2x8 / code1x8 x {matmat, matmat_dequant, matmat_dequant_transposed}), bound to liblutvq instead of the
JIT-built `codebook_cuda` extension.  Anything that reaches the kernels through `torch.ops.aqlm` (the HF AQLM integration,
captured graphs as in notebooks/aqlm_cuda_graph.ipynb) binds to these.  CUDA only: there is no CPU implementation to
register, so a CPU tensor raises NotImplementedError from the dispatcher.

The schema takes `Tensor? bias` (the reference writes `Tensor bias` but passes None for bias-free layers)."""
from __future__ import annotations

import torch

from . import ops

_SCHEMA = "(Tensor input, Tensor codes, Tensor codebooks, Tensor scales, Tensor? bias=None) -> Tensor"
_FORWARD = {
    "code1x16_matmat": ops.code1x16_matmat,
    "code1x16_matmat_dequant": ops.code1x16_matmat_dequant,
    "code2x8_matmat": ops.code2x8_matmat,
    "code2x8_matmat_dequant": ops.code2x8_matmat_dequant,
    "code1x8_matmat": ops.code1x8_matmat,
    "code1x8_matmat_dequant": ops.code1x8_matmat_dequant,
    "codeKx8_matmat": ops.codeKx8_matmat,                  # not a reference op name: N in {1..8} x 256 (cuda_kernel.cpp:521-548 codeKx8 family)
}
_TRANSPOSED = {
    "code1x16_matmat_dequant_transposed": ops.code1x16_matmat_dequant_transposed,
    "code2x8_matmat_dequant_transposed": ops.code2x8_matmat_dequant_transposed,
    "code1x8_matmat_dequant_transposed": ops.code1x8_matmat_dequant_transposed,
}
_registered = False
_lib_handle = None


def _meta_forward(input, codes, codebooks, scales, bias=None):
    return torch.empty(input.shape[:-1] + (codes.shape[0],), device=input.device, dtype=input.dtype)           # cuda_kernel.py:21-23


def _meta_transposed(input, codes, codebooks, scales, bias=None):
    return torch.empty(input.shape[:-1] + (codes.shape[1] * codebooks.shape[3],), device=input.device, dtype=input.dtype)   # :48-52


def register() -> None:
    """idempotent; called on import of rvv_lut_vq_b200.linear / kernel_selector"""
    global _registered, _lib_handle
    if _registered:
        return
    lib = torch.library.Library("aqlm", "FRAGMENT")
    for table, meta in ((_FORWARD, _meta_forward), (_TRANSPOSED, _meta_transposed)):
        for name, fn in table.items():
            lib.define(name + _SCHEMA)
            lib.impl(name, (lambda f: lambda input, codes, codebooks, scales, bias=None: f(input, codes, codebooks, scales, bias))(fn), "CUDA")
            torch.library.register_fake(f"aqlm::{name}", meta, lib=lib)
    _lib_handle = lib
    _registered = True


register()
