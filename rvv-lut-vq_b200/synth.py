"""Synthetic AQLM-2x8 layers with the value distributions of the reference's own benchmarks
(reference: benchmark/kernel/rq_lut/benchmark_rq_gemm_gpu.py:83-97, vq_dataflow_sim/vq.py:52,64):
x ~ N(0,1), codes ~ U{0..255} stored as int8, codebooks ~ N(0,1), scales = 1 (GPU script) or N(0,1) (CPU script)."""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np

# Llama-2 shapes (in_features, out_features): SURVEY 3.4 / 9
LLAMA2_7B_LAYER = {"q": (4096, 4096), "k": (4096, 4096), "v": (4096, 4096), "o": (4096, 4096),
                   "gate": (4096, 11008), "up": (4096, 11008), "down": (11008, 4096)}
LLAMA2_70B_LAYER = {"q": (8192, 8192), "k": (8192, 1024), "v": (8192, 1024), "o": (8192, 8192),
                    "gate": (8192, 28672), "up": (8192, 28672), "down": (28672, 8192)}
SWEEP_7B = [(4096, 4096), (4096, 11008), (11008, 4096)]          # BASELINE.json configs[1]


def gemv_algorithmic_bytes(in_features: int, out_features: int, num_codebooks: int = 2) -> int:
    """O*(I/8)*N codes + N*256*8*2 codebooks + 2I x + 2O scales + 2O y"""
    return out_features * (in_features // 8) * num_codebooks + num_codebooks * 256 * 8 * 2 + 2 * in_features + 4 * out_features


def make_layer_np(in_features: int, out_features: int, num_codebooks: int = 2, seed: int = 0, scales: str = "randn",
                  bias: bool = False, batch: int = 1) -> Dict[str, np.ndarray]:
    """numpy (CPU) generator for the parity tests: fp16-representable values so fp16 GPU runs see identical operands."""
    rng = np.random.default_rng(seed)
    G = in_features // 8
    idx = rng.integers(0, 256, size=(out_features, G, num_codebooks), dtype=np.uint8)
    out = {
        "x": rng.standard_normal((batch, in_features)).astype(np.float16),
        "idx": idx,
        "codes": idx.view(np.int8),                                   # pack_int_data: same bits, signed view
        "codebooks": rng.standard_normal((num_codebooks, 256, 1, 8)).astype(np.float16),
        "scales": (rng.standard_normal((out_features, 1, 1, 1)) if scales == "randn"
                   else np.ones((out_features, 1, 1, 1))).astype(np.float16),
        "bias": rng.standard_normal(out_features).astype(np.float16) if bias else None,
    }
    return out


def make_layer_torch(in_features: int, out_features: int, device, dtype=None, num_codebooks: int = 2, seed: int = 0,
                     scales: str = "ones", bias: bool = False):
    """device-side generator for benchmarks (full Llama shapes; mirrors benchmark_rq_gemm_gpu.py:83-97)."""
    import torch
    dtype = dtype or torch.float16
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    G = in_features // 8
    codes = torch.randint(0, 256, (out_features, G, num_codebooks), device=device, generator=g, dtype=torch.int16)
    codes = codes.to(torch.uint8).view(torch.int8)                    # pack_int_data semantics (two's complement view)
    codebooks = torch.randn((num_codebooks, 256, 1, 8), device=device, generator=g, dtype=torch.float32).to(dtype)
    if scales == "ones":
        sc = torch.ones((out_features, 1, 1, 1), device=device, dtype=dtype)
    else:
        sc = torch.randn((out_features, 1, 1, 1), device=device, generator=g, dtype=torch.float32).to(dtype)
    b = torch.randn(out_features, device=device, generator=g, dtype=torch.float32).to(dtype) if bias else None
    return {"codes": codes, "codebooks": codebooks, "scales": sc, "bias": b}
