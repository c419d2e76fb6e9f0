"""Code packing helpers with the reference's names and semantics
(reference: benchmark/AQLM/inference_lib/src/aqlm/utils.py:11-31)."""
from __future__ import annotations

import torch


def get_int_dtype(nbits: int) -> torch.dtype:
    """utils.py:11-20"""
    if nbits <= 8:
        return torch.int8
    if nbits <= 16:
        return torch.int16
    if nbits <= 32:
        return torch.int32
    if nbits <= 64:
        return torch.int64
    raise ValueError(f"No dtype available for {nbits}-bit codebooks")


@torch.inference_mode()
def pack_int_data(data: torch.Tensor, nbits: int) -> torch.Tensor:
    """utils.py:23-26: unsigned indices stored in the signed dtype of the same width.
    Like the reference this wraps values >= 2^(nbits-1) IN PLACE before casting."""
    data[data >= 2 ** (nbits - 1)] -= 2**nbits
    return data.to(get_int_dtype(nbits))


@torch.inference_mode()
def unpack_int_data(data: torch.Tensor, nbits: int) -> torch.Tensor:
    """utils.py:29-31"""
    return data.to(torch.int64) % (2**nbits)
