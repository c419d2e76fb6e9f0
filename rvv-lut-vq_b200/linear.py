"""QuantizedLinear with the reference's constructor, parameter names/shapes, lazy op selection and batch-size dispatch
rule (reference: benchmark/AQLM/inference_lib/src/aqlm/inference.py:11-142, kernel_selector.py:21-145).

Formats with a kernel: AQLM 2x8 (the north-star hot path: fused LUT-GEMV on a prepacked code image for batch <= 6,
dequantize-into-tensor-memory + tcgen05 GEMM above), 1x8 / Kx8 (N <= 8 codebooks of 256), 1x16 (65536-entry codebook, groups of 8 or
16).  CUDA only; anything else raises NotImplementedError instead of taking a generic / Triton / CPU route."""
from __future__ import annotations

import math
from typing import Any, Optional

import torch
import torch.nn as nn

from . import ops
from .kernel_selector import get_backward_pass_kernel, get_forward_pass_kernel
from .utils import get_int_dtype

GEMV_BATCH_THRESHOLD = 6  # inference.py:95: prod(input.shape[:-1]) <= 6 -> gemv op, else gemm op
# (The reference's rule, kept as the default.  Measured on B200 at 4096x11008 the two ops cross at 3-4 rows: grouped GEMV 11.8 / 16.3 / 19.6 us
# at 3 / 4 / 6 rows, the tcgen05 prefill kernel 14.1 us for any batch up to 16 -- DESIGN 4.1 / 4.4; lower this constant to 3 to take it.)


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
        # MATMUL_OPS (inference.py:63-66), selected lazily on the first forward
        self.gemv_op = None
        self.gemm_op = None
        self._packed = None          # 2x8 GEMV code image (the analogue of the CPU code permutation, inference.py:74-83)
        self._packed_key = None      # (codes.data_ptr(), codes._version, device): a reload / in-place edit / .to() rebuilds it
        # X_READY contract of include/lutvq.h: set True only when this layer's input was COMPLETE before the whole run of
        # overlapped launches began (a resident buffer) -- under programmatic dependent launch "not produced by the kernel
        # right before" is NOT enough (producer -> A -> B(X_READY) can read x before the producer finished).
        self.x_ready = False

    def use_gemv_rule(self, input: torch.Tensor) -> bool:
        return math.prod(input.shape[:-1]) <= GEMV_BATCH_THRESHOLD

    def prepare_matmul_op(self, input: torch.Tensor) -> None:
        """inference.py:74-93 without the CPU branch"""
        if not input.is_cuda:
            raise NotImplementedError("lutvq has no CPU path: move the module and its input to a CUDA device")
        self.gemv_op = _get_autograd_matmul_op(get_forward_pass_kernel(self.codebooks, False), get_backward_pass_kernel(self.codebooks, False))
        self.gemm_op = _get_autograd_matmul_op(get_forward_pass_kernel(self.codebooks, True), get_backward_pass_kernel(self.codebooks, True))

    def _is_2x8(self) -> bool:
        return (self.num_codebooks, self.codebook_size, self.out_group_size, self.in_group_size) == (2, 256, 1, 8)

    def _packed_image(self, input: torch.Tensor) -> torch.Tensor:
        key = (self.codes.data_ptr(), self.codes._version, self.codes.device)
        if self._packed is None or self._packed_key != key:
            self._packed = ops.code2x8_prepack(self.codes)
            self._packed_key = key
        return self._packed

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        if self.gemv_op is None:
            self.prepare_matmul_op(input)
        if not input.is_cuda or input.device != self.codes.device:
            raise NotImplementedError(f"input on {input.device}, module on {self.codes.device}: lutvq has no CPU path and does not move tensors")
        if self.use_gemv_rule(input):
            if self._is_2x8() and not (torch.is_grad_enabled() and input.requires_grad):
                # the hot path: fused LUT-GEMV on the prepacked image (bit-identical to torch.ops.aqlm.code2x8_matmat)
                return ops.code2x8_matmat_packed(input, self._packed_image(input), self.out_features, self.codebooks, self.scales,
                                                 self.bias, x_ready=self.x_ready)
            return self.gemv_op.apply(input, self.codes, self.codebooks, self.scales, self.bias)
        return self.gemm_op.apply(input, self.codes, self.codebooks, self.scales, self.bias)


def _get_autograd_matmul_op(forward_pass_kernel, backward_pass_kernel):
    """inference.py:99-142: forward through the selected op, grad_input through the transposed op."""

    class _QuantizedMatmul(torch.autograd.Function):
        @staticmethod
        def forward(ctx: Any, input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                    bias: Optional[torch.Tensor]) -> torch.Tensor:
            ctx.save_for_backward(input, codes, codebooks, scales, bias)
            return forward_pass_kernel(input, codes, codebooks, scales, bias)

        @staticmethod
        def backward(ctx, grad_output: torch.Tensor):
            input, codes, codebooks, scales, bias = ctx.saved_tensors
            return (backward_pass_kernel(grad_output, codes, codebooks, scales, None), None, None, None, None)

    return _QuantizedMatmul
