"""AQLM checkpoints in the HF layout: what ``benchmark/AQLM/convert_to_hf.py:50-100,164`` writes.

Layout (convert_to_hf.py:50-79): ``pytorch_model.bin`` (or ``*.safetensors``) holds, for every quantized linear ``<prefix>``
(e.g. ``model.layers.3.self_attn.q_proj``), the QuantizedLinear state ``<prefix>.codes`` (packed signed ints holding unsigned
indices, pack_int_data), ``<prefix>.codebooks`` [N, 2**nbits, og, ig] fp16, ``<prefix>.scales`` [O/og, 1, 1, 1] fp16 and
optionally ``<prefix>.bias``; everything else (embeddings, norms, lm_head, the layers named in ``linear_weights_not_to_quantize``)
as plain fp16 tensors.  ``config.json`` carries ``quantization_config = {quant_method: "aqlm", nbits_per_codebook, num_codebooks,
out_group_size, in_group_size, linear_weights_not_to_quantize}`` (convert_to_hf.py:82-100).

This module reads that layout WITHOUT transformers / accelerate / the hub (none of them works offline in this image,
SURVEY 9): ``load_checkpoint`` -> (config dict, state dict), ``quantized_linears`` -> {prefix: QuantizedLinear} on the GPU, and
``load_llama_decoder`` fills a ``decode.Decoder`` (Llama-2 shapes, AQLM 2x8) from a checkpoint directory, so the decode
benchmark can run real ``ISTA-DASLab/*-AQLM-2Bit-2x8`` weights when they are on disk.  ``save_checkpoint`` writes the layout
(used by the tests to round-trip; the reference's converter is the real producer)."""
from __future__ import annotations

import glob
import json
import os
from typing import Dict, Optional, Tuple

import torch

QUANT_SUFFIXES = ("codes", "codebooks", "scales")


def load_checkpoint(path: str) -> Tuple[dict, Dict[str, torch.Tensor]]:
    """-> (config.json as a dict, flat state dict on the CPU).  Reads pytorch_model*.bin and/or *.safetensors."""
    with open(os.path.join(path, "config.json")) as f:
        config = json.load(f)
    q = config.get("quantization_config")
    if not q or q.get("quant_method") != "aqlm":
        raise ValueError(f"{path}/config.json has no AQLM quantization_config (convert_to_hf.py:82-100)")
    state: Dict[str, torch.Tensor] = {}
    for fn in sorted(glob.glob(os.path.join(path, "pytorch_model*.bin"))):
        state.update(torch.load(fn, map_location="cpu", weights_only=True))
    st_files = sorted(glob.glob(os.path.join(path, "*.safetensors")))
    if st_files:
        try:
            from safetensors.torch import load_file
        except ImportError as e:  # pragma: no cover
            raise ImportError("the checkpoint is in safetensors format and the safetensors package is not installed") from e
        for fn in st_files:
            state.update(load_file(fn, device="cpu"))
    if not state:
        raise FileNotFoundError(f"no pytorch_model*.bin / *.safetensors under {path}")
    return config, state


def save_checkpoint(path: str, config: dict, state: Dict[str, torch.Tensor]) -> None:
    """the layout of convert_to_hf.py:164-170 (pytorch_model.bin + config.json)"""
    os.makedirs(path, exist_ok=True)
    torch.save({k: v.detach().cpu() for k, v in state.items()}, os.path.join(path, "pytorch_model.bin"))
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(config, f, indent=4)


def quantized_prefixes(state: Dict[str, torch.Tensor]):
    """every <prefix> that has <prefix>.codes / .codebooks / .scales"""
    pre = sorted({k[: -len(".codes")] for k in state if k.endswith(".codes")})
    for p in pre:
        for s in QUANT_SUFFIXES:
            if f"{p}.{s}" not in state:
                raise KeyError(f"{p}.{s} missing: an AQLM linear needs codes, codebooks and scales (inference.py:39-61)")
    return pre


def quantized_linears(config: dict, state: Dict[str, torch.Tensor], device, dtype=torch.float16, prefix_filter: str = ""):
    """{prefix: QuantizedLinear on `device`} for every quantized linear of the checkpoint, shapes checked against
    quantization_config (nbits_per_codebook, num_codebooks, out_group_size, in_group_size)."""
    from .linear import QuantizedLinear
    from .utils import get_int_dtype
    q = config["quantization_config"]
    nbits, N, og, ig = q["nbits_per_codebook"], q["num_codebooks"], q["out_group_size"], q["in_group_size"]
    out = {}
    for p in quantized_prefixes(state):
        if prefix_filter and not p.startswith(prefix_filter):
            continue
        codes, cb, sc = state[p + ".codes"], state[p + ".codebooks"], state[p + ".scales"]
        if tuple(cb.shape) != (N, 2 ** nbits, og, ig) or codes.dtype != get_int_dtype(nbits) or codes.shape[2] != N:
            raise ValueError(f"{p}: tensors do not match quantization_config {q} (codebooks {tuple(cb.shape)}, codes {tuple(codes.shape)} {codes.dtype})")
        O, I = codes.shape[0] * og, codes.shape[1] * ig
        bias = state.get(p + ".bias")
        m = QuantizedLinear(I, O, ig, og, N, nbits, bias=bias is not None, device=device, dtype=dtype)
        sd = {"codes": codes, "codebooks": cb.to(dtype), "scales": sc.to(dtype)}
        if bias is not None:
            sd["bias"] = bias.to(dtype)
        m.load_state_dict(sd)
        out[p] = m
    return out


_LLAMA_LINEARS = {"q": "self_attn.q_proj", "k": "self_attn.k_proj", "v": "self_attn.v_proj", "o": "self_attn.o_proj",
                  "gate": "mlp.gate_proj", "up": "mlp.up_proj", "down": "mlp.down_proj"}


def llama_decoder_config(config: dict):
    """config.json (LlamaConfig fields) -> decode.DecoderConfig"""
    from .decode import DecoderConfig
    H, heads = config["hidden_size"], config["num_attention_heads"]
    return DecoderConfig(config.get("_name_or_path", "llama-aqlm") or "llama-aqlm", H, config["intermediate_size"], config["num_hidden_layers"], heads,
                         config.get("num_key_value_heads", heads), vocab=config["vocab_size"], head_dim=H // heads,
                         rope_theta=float(config.get("rope_theta", 10000.0)), eps=float(config.get("rms_norm_eps", 1e-5)))


def load_llama_decoder(path: str, device, world: int = 1, rank: int = 0, group=None, max_seq: int = 512, exchange: str = "ll"):
    """A decode.Decoder whose weights come from an AQLM 2x8 Llama checkpoint directory (every rank keeps its output-channel
    shard of every linear).  Raises for other formats: the decode loop is built on the 2x8 LUT-GEMV."""
    from . import ops
    from .decode import Decoder
    from .sharded import shard_rows
    config, state = load_checkpoint(path)
    q = config["quantization_config"]
    if (q["num_codebooks"], q["nbits_per_codebook"], q["out_group_size"], q["in_group_size"]) != (2, 8, 1, 8):
        raise NotImplementedError(f"the decode loop takes AQLM 2x8 checkpoints, got {q}")
    cfg = llama_decoder_config(config)
    dec = Decoder(cfg, device, world, rank, group, max_seq=max_seq, exchange=exchange if (world > 1 or group is not None) else "push", skeleton=True)
    dev = dec.dev
    f16 = lambda t: t.to(device=dev, dtype=torch.float16)
    for li, L in enumerate(dec.layers):
        base = f"model.layers.{li}."
        for name, hf in _LLAMA_LINEARS.items():
            lin = L[name]
            p = base + hf
            codes = state[p + ".codes"]
            assert tuple(codes.shape) == (lin.out_f, lin.in_f // 8, 2), f"{p}.codes {tuple(codes.shape)} vs ({lin.out_f}, {lin.in_f // 8}, 2)"
            b, e = shard_rows(lin.out_f, world, rank)
            lin.packed = ops.code2x8_prepack(codes[b:e].contiguous().to(dev))
            lin.codebooks = f16(state[p + ".codebooks"]).contiguous()
            lin.scales = f16(state[p + ".scales"]).reshape(-1)[b:e].contiguous()
        L["norm1"] = f16(state[base + "input_layernorm.weight"])
        L["norm2"] = f16(state[base + "post_attention_layernorm.weight"])
    dec.embed = f16(state["model.embed_tokens.weight"])
    dec.final_norm = f16(state["model.norm.weight"])
    dec.lm_head = f16(state.get("lm_head.weight", state["model.embed_tokens.weight"]))
    return dec
