"""QuantizedLinear with the reference's constructor, parameter names/shapes and batch-size dispatch rule
(reference: benchmark/AQLM/inference_lib/src/aqlm/inference.py:11-96, kernel_selector.py:21-102).

Only the AQLM 2x8 format (2 codebooks x 256 entries, in_group_size 8, out_group_size 1) on CUDA is
implemented -- the north-star hot path.  Anything else raises NotImplementedError instead of silently
taking a generic/Triton/CPU route (the reference's multi-backend dispatch is deliberately not reproduced)."""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn

from . import ops
from .utils import get_int_dtype

GEMV_BATCH_THRESHOLD = 6  # inference.py:95: prod(input.shape[:-1]) <= 6 -> gemv op, else gemm op


class QuantizedLinear(nn.Module):
    def __init__(self, in_features: int, out_features: int, in_group_size: int, out_group_size: int,
                 num_codebooks: int, nbits_per_codebook: int, bias: bool = True, device=None, dtype=None):
        factory_kwargs = {"device": device, "dtype": dtype}
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        assert in_features % in_group_size == 0
        assert out_features % out_group_size == 0
        num_out_groups = out_features // out_group_size
        num_in_groups = in_features // in_group_size
        self.out_group_size, self.in_group_size = out_group_size, in_group_size
        self.num_codebooks = num_codebooks
        self.nbits_per_codebook = nbits_per_codebook
        self.codebook_size = 2**nbits_per_codebook
        # same parameter names / shapes / dtypes as inference.py:39-61 so state dicts are interchangeable
        self.codebooks = nn.Parameter(
            torch.empty((num_codebooks, self.codebook_size, out_group_size, in_group_size), **factory_kwargs),
            requires_grad=False)
        self.codes = nn.Parameter(
            torch.empty((num_out_groups, num_in_groups, num_codebooks), device=device,
                        dtype=get_int_dtype(nbits_per_codebook)), requires_grad=False)
        self.scales = nn.Parameter(torch.empty((num_out_groups, 1, 1, 1), **factory_kwargs), requires_grad=False)
        if bias:
            self.bias = nn.Parameter(torch.empty(out_features, **factory_kwargs), requires_grad=False)
        else:
            self.register_parameter("bias", None)
        self._packed = None          # GEMV code image, built on first use (prepare_matmul_op analogue, inference.py:74-83)
        self.x_ready = False         # set True when this layer's input is never the output of the kernel right before it

    def _check_format(self, input: torch.Tensor) -> None:
        fmt = (self.num_codebooks, self.codebook_size, self.out_group_size, self.in_group_size)
        if fmt != (2, 256, 1, 8):
            raise NotImplementedError(
                f"lutvq implements the AQLM 2x8 format (2 codebooks x 256, groups 1x8); got {fmt}")
        if not input.is_cuda:
            raise NotImplementedError("lutvq has no CPU path: move the module and its input to a CUDA device")

    def use_gemv_rule(self, input: torch.Tensor) -> bool:
        return math.prod(input.shape[:-1]) <= GEMV_BATCH_THRESHOLD

    @torch.no_grad()
    def forward(self, input: torch.Tensor) -> torch.Tensor:
        self._check_format(input)
        if self.use_gemv_rule(input):
            if self._packed is None or self._packed.device != input.device:
                self._packed = ops.code2x8_prepack(self.codes)
            return ops.code2x8_matmat_packed(input, self._packed, self.out_features, self.codebooks, self.scales,
                                             self.bias, x_ready=self.x_ready)
        return ops.code2x8_matmat_dequant(input, self.codes, self.codebooks, self.scales, self.bias)
