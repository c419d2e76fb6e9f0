"""Torch-tensor front end of liblutvq.so with the reference's operator names, argument order and error
behaviour (reference: benchmark/AQLM/inference_lib/src/aqlm/inference_kernels/cuda_kernel.cpp:356-484 and
the torch.library schema `(Tensor input, Tensor codes, Tensor codebooks, Tensor scales, Tensor bias) -> Tensor`
in cuda_kernel.py:13-132).  Torch is plumbing only: tensors supply device pointers and the current stream;
every op is one or more launches of our own sm_100a kernels through the C-ABI.  No CPU path exists."""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib
from ._lib import check, lib

_DT = {torch.float16: _lib.LUTVQ_F16, torch.bfloat16: _lib.LUTVQ_BF16}
_workspaces: Dict[Tuple[int, int], torch.Tensor] = {}


def check_use_bfloat16(t: torch.Tensor) -> int:
    """cuda_kernel.cpp:9-25: only fp16 / bf16; anything else is NotImplementedError."""
    if t.dtype not in _DT:
        raise NotImplementedError(
            f"AQLM CUDA kernels only support float16 and bfloat16. Got {t.dtype}. "
            "Please specify the correct `torch_dtype` when loading the model.")
    return _DT[t.dtype]


def _require_cuda(*tensors: Optional[torch.Tensor]) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("lutvq ops run on CUDA tensors only (there is no CPU fallback)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"tensors on different devices: {dev} vs {t.device}")
    return dev


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream(dev: torch.device) -> ctypes.c_void_p:
    return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


XREADY_WS_RING = 8          # include/lutvq.h LUTVQ_XREADY_WS_RING
_ws_ring_pos: Dict[Tuple[int, int], int] = {}


_gemm_workspaces: Dict[Tuple[int, int], torch.Tensor] = {}
_stage_workspaces: Dict[Tuple[int, int], torch.Tensor] = {}


def _key(dev: torch.device) -> Tuple[int, int]:
    return (dev.index if dev.index is not None else torch.cuda.current_device(), torch.cuda.current_stream(dev).cuda_stream)


def gemm_workspace(dev: torch.device, prob_m: int, prob_k: int, batch: int) -> torch.Tensor:
    """Scratch of the prefill GEMM (split-k partial accumulators; can be tens of MB), cached per (device, stream)."""
    need = int(lib().lutvq_workspace_bytes_ex(prob_m, prob_k, batch))
    key = _key(dev)
    ws = _gemm_workspaces.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.zeros(max(need, 1 << 20), dtype=torch.uint8, device=dev)
        _gemm_workspaces[key] = ws
    return ws


def stage_workspace(dev: torch.device, prob_m: int, prob_k: int) -> torch.Tensor:
    """Scratch of the reference-layout GEMV entry points: split-k words + the staging area for the re-laid codes
    (lutvq_workspace_bytes), cached per (device, stream).  Callers that prepack once never need it."""
    need = int(lib().lutvq_workspace_bytes(prob_m, prob_k))
    key = _key(dev)
    ws = _stage_workspaces.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.zeros(need, dtype=torch.uint8, device=dev)
        _stage_workspaces[key] = ws
    return ws


def workspace(dev: torch.device, prob_m: int, prob_k: int, rotate: bool = False, stream: Optional[int] = None) -> torch.Tensor:
    """Zero-initialised split-k words for the packed entry points, cached per (device, stream) as include/lutvq.h
    requires.  The cache holds a ring of XREADY_WS_RING regions; plain calls always use region 0, ``rotate=True``
    (LUTVQ_FLAG_X_READY calls) steps through regions 1.. so that overlapping launches never share a region and never
    collide with a plain call's region 0.  (Inside a CUDA graph the region is baked at capture.)"""
    need = _ws_need.get((prob_m, prob_k))
    if need is None:
        need = _ws_need[(prob_m, prob_k)] = (int(lib().lutvq_packed_workspace_bytes(prob_m, prob_k)) + 255) // 256 * 256
    key = _key(dev) if stream is None else (dev.index, stream)
    ent = _ws_regions.get(key)
    if ent is None or ent[1] < need:
        ws = torch.zeros(max(need, 1 << 17) * XREADY_WS_RING, dtype=torch.uint8, device=dev)
        _workspaces[key] = ws
        slot = ws.numel() // XREADY_WS_RING
        ent = ([ws[i * slot:(i + 1) * slot] for i in range(XREADY_WS_RING)], slot)
        _ws_regions[key] = ent
    if not rotate:
        return ent[0][0]
    i = _ws_ring_pos.get(key, 0) % (XREADY_WS_RING - 1) + 1
    _ws_ring_pos[key] = i
    return ent[0][i]


_batch_workspaces: Dict[Tuple[int, int], torch.Tensor] = {}
BATCH_GROUPED = os.environ.get("LUTVQ_BATCH_GROUPED", "1") != "0"      # experiment switch: 0 = one launch per row of a multi-row call


def batch_workspace(dev: torch.device, prob_m: int, prob_k: int, rows: int, stream: Optional[int] = None) -> torch.Tensor:
    """min(rows, XREADY_WS_RING) zero-initialised regions for ONE multi-row GEMV call (include/lutvq.h, lutvq_code2x8_matmat): up to 8
    rows of the call then go out as one launch.  A buffer of its own per (device, stream), so it never collides with the X_READY
    ring of `workspace()`; the next multi-row call on the stream reuses it (it waits for everything before it)."""
    need = _ws_need.get((prob_m, prob_k))
    if need is None:
        need = _ws_need[(prob_m, prob_k)] = (int(lib().lutvq_packed_workspace_bytes(prob_m, prob_k)) + 255) // 256 * 256
    total = need * min(rows, XREADY_WS_RING)
    key = _key(dev) if stream is None else (dev.index, stream)
    ws = _batch_workspaces.get(key)
    if ws is None or ws.numel() < total:
        ws = torch.zeros(max(total, 1 << 20), dtype=torch.uint8, device=dev)
        _batch_workspaces[key] = ws
    return ws[:total]


_ws_regions: Dict[Tuple[int, int], tuple] = {}
_ws_need: Dict[Tuple[int, int], int] = {}


def workspace_zero_bytes(prob_m: int) -> int:
    """the part of a packed workspace every call hands back all-zero: 256-byte header + one arrival counter per 32-row set"""
    return 256 + (4 * ((prob_m + 31) // 32) + 255) // 256 * 256


def _flat_scales(scales: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    return None if scales is None else scales.reshape(-1).contiguous()


def code2x8_matvec(A: torch.Tensor, B: torch.Tensor, C: torch.Tensor, codebook: torch.Tensor, use_bfloat16: bool) -> None:
    """cuda_kernel.cpp:356-385: C[prob_m] = W(A) . B[prob_k], written into the caller's C."""
    dev = _require_cuda(A, B, C, codebook)
    prob_m, prob_k = C.size(0), B.size(0)
    with torch.cuda.device(dev):
        ws = stage_workspace(dev, prob_m, prob_k)
        check(lib().lutvq_code2x8_matvec(_ptr(A), _ptr(B), _ptr(C), _ptr(codebook), prob_m, prob_k,
                                         _lib.LUTVQ_BF16 if use_bfloat16 else _lib.LUTVQ_F16,
                                         _ptr(ws), ws.numel(), _stream(dev)))


def _matmat(fn_name: str, input, codes, codebooks, scales, bias, flags: Optional[int] = None) -> torch.Tensor:
    dtype = check_use_bfloat16(input)
    dev = _require_cuda(input, codes, codebooks, scales, bias)
    out_features = codes.size(0) * codebooks.size(2)
    in_features = input.size(-1)
    flat_input = input.reshape(-1, in_features).contiguous()
    batch = flat_input.size(0)
    flat_output = torch.empty((batch, out_features), dtype=input.dtype, device=dev)
    codes_c, cb_c = codes.contiguous(), codebooks.contiguous()
    sc, bi = _flat_scales(scales), (None if bias is None else bias.contiguous())
    with torch.cuda.device(dev):
        if fn_name == "lutvq_code2x8_matmat_dequant":
            ws = gemm_workspace(dev, out_features, in_features, batch)
        elif batch > 1 and BATCH_GROUPED and not (flags or 0) & _lib.FLAG_X_READY:
            ws = batch_workspace(dev, out_features, in_features, batch)
        else:
            ws = stage_workspace(dev, out_features, in_features)
        fn = getattr(lib(), fn_name)
        if flags is None:
            rc = fn(_ptr(flat_input), _ptr(codes_c), _ptr(cb_c), _ptr(sc), _ptr(bi), _ptr(flat_output), batch,
                    out_features, in_features, dtype, _ptr(ws), ws.numel(), _stream(dev))
        else:
            rc = fn(_ptr(flat_input), _ptr(codes_c), _ptr(cb_c), _ptr(sc), _ptr(bi), _ptr(flat_output), batch,
                    out_features, in_features, dtype, flags, _ptr(ws), ws.numel(), _stream(dev))
        check(rc)
    return flat_output.reshape(tuple(input.shape[:-1]) + (out_features,))


def code2x8_matmat(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                   bias: Optional[torch.Tensor] = None, *, lut_fp16: bool = False, x_ready: bool = False) -> torch.Tensor:
    """cuda_kernel.cpp:387-421 (+ epilogue :95-111): the decode / small-batch path (fused LUT-GEMV per row).
    ``x_ready``: LUTVQ_FLAG_X_READY -- the caller's promise of include/lutvq.h: the input was COMPLETE before the whole run of
    back-to-back launches this call belongs to started ("not produced by the kernel right before" is NOT enough: with
    producer -> A -> B(x_ready), B can read x before the producer has finished), and the workspace ring is not shared."""
    flags = (_lib.FLAG_LUT_FP16 if lut_fp16 else 0) | (_lib.FLAG_X_READY if x_ready else 0)
    if flags:
        return _matmat("lutvq_code2x8_matmat_ex", input, codes, codebooks, scales, bias, flags)
    return _matmat("lutvq_code2x8_matmat", input, codes, codebooks, scales, bias)


def code2x8_prepack(codes: torch.Tensor) -> torch.Tensor:
    """One-off re-layout of QuantizedLinear.codes for the GEMV (the counterpart of the CPU permutation in
    inference.py:78-83); returns an opaque uint8 image for code2x8_matmat_packed."""
    dev = _require_cuda(codes)
    prob_m, prob_k = codes.size(0), codes.size(1) * 8
    packed = torch.empty(int(lib().lutvq_code2x8_packed_bytes(prob_m, prob_k)), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_code2x8_prepack(_ptr(codes.contiguous()), _ptr(packed), prob_m, prob_k, _stream(dev)))
    return packed


def code2x8_matmat_packed(input: torch.Tensor, packed: torch.Tensor, out_features: int, codebooks: torch.Tensor,
                          scales: torch.Tensor, bias: Optional[torch.Tensor] = None, *, lut_fp16: bool = False,
                          x_ready: bool = False) -> torch.Tensor:
    """code2x8_matmat on a prepacked code image (bit-identical results).  This is QuantizedLinear's per-token call, so the
    host side is kept lean: no device context switch when the tensors already live on the current device, raw integer
    pointers into the ctypes call, one dict lookup for the cached workspace."""
    dtype = _DT.get(input.dtype)
    if dtype is None:
        check_use_bfloat16(input)
    if not (input.is_cuda and packed.is_cuda and codebooks.is_cuda):
        raise RuntimeError("lutvq ops run on CUDA tensors only (there is no CPU fallback)")
    dev = input.device
    if packed.device != dev or codebooks.device != dev:
        _require_cuda(input, packed, codebooks, scales, bias)
    in_features = input.shape[-1]
    flat_input = input if input.is_contiguous() else input.contiguous()
    batch = flat_input.numel() // in_features
    flat_output = torch.empty((batch, out_features), dtype=input.dtype, device=dev)
    flags = (_lib.FLAG_LUT_FP16 if lut_fp16 else 0) | (_lib.FLAG_X_READY if x_ready else 0)
    sc = None if scales is None else (scales if scales.is_contiguous() else scales.contiguous())
    bi = None if bias is None else (bias if bias.is_contiguous() else bias.contiguous())
    cb = codebooks if codebooks.is_contiguous() else codebooks.contiguous()
    prev = _cuda_get_device()
    switch = prev != dev.index
    if switch:
        torch.cuda.set_device(dev)
    try:
        stream = _cuda_raw_stream(dev.index)
        if batch > 1 and BATCH_GROUPED and not x_ready:
            ws = batch_workspace(dev, out_features, in_features, batch, stream)
        else:
            ws = workspace(dev, out_features, in_features, rotate=x_ready, stream=stream)
        rc = _fn_matmat_packed()(flat_input.data_ptr(), packed.data_ptr(), cb.data_ptr(), 0 if sc is None else sc.data_ptr(),
                                 0 if bi is None else bi.data_ptr(), flat_output.data_ptr(), batch, out_features, in_features, dtype, flags,
                                 ws.data_ptr(), ws.numel(), stream)
        if rc != 0:
            check(rc)
    finally:
        if switch:
            torch.cuda.set_device(prev)
    return flat_output if input.dim() == 2 else flat_output.reshape(tuple(input.shape[:-1]) + (out_features,))


# the raw bindings behind torch.cuda.current_device() / current_stream().cuda_stream (each saves 1-2 us of Python per call)
_cuda_get_device = getattr(torch._C, "_cuda_getDevice", None) or torch.cuda.current_device
_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _cuda_raw_stream(index: int) -> int:
    return _raw_stream(index) if _raw_stream is not None else torch.cuda.current_stream(index).cuda_stream


_matmat_packed_fn = None


def _fn_matmat_packed():
    global _matmat_packed_fn
    if _matmat_packed_fn is None:
        _matmat_packed_fn = lib().lutvq_code2x8_matmat_packed
    return _matmat_packed_fn


def code2x8_matmat_dequant_fused(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                                 bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """cuda_kernel.cpp:450-484 as ONE kernel: dequantize into tensor memory + tcgen05 GEMM (W never exists in HBM), the
    reference's epilogue order (round(acc) * scale + bias).  Batches above 256 rows are chunked (the codes are re-read per chunk)."""
    return _matmat("lutvq_code2x8_matmat_dequant", input, codes, codebooks, scales, bias)


def code2x8_matmat_dequant_two_step(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                                    bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """cuda_kernel.cpp:450-484 the way the reference does it: dequantize W once (our kernel, scale folded in as ONE rounded
    multiply per weight like code2x8_dequant, cuda_kernel.cpp:445), then the plain library GEMM the reference also calls
    (F.linear = cuBLAS, bias in its epilogue).  W is written once and read from L2; the cost of dequantizing is paid once
    whatever the batch is."""
    check_use_bfloat16(input)
    W = code2x8_dequant(codes, codebooks, scales)
    flat = input.reshape(-1, input.size(-1))
    return torch.nn.functional.linear(flat, W, bias).reshape(tuple(input.shape[:-1]) + (W.size(0),))


# rows from which dequantize-once + library GEMM beats the fused kernel.  Measured on B200, 4096x11008, a ring of 24 L2-cold
# layers, scales applied (bench.py `prefill`, profiles/r02_bench_line.json): fused 14.0 / 14.4 / 15.0 / 18.0 / 27.9 us at
# 16 / 32 / 64 / 128 / 256 rows, then 58 us at 512 and ~250 us at 2048 (one launch takes <= 256 rows: chunks re-read the codes);
# two-step 38.4 / 38.0 / 38.4 / 35.9 / 41.9 us at 16 .. 256 rows, 58 at 512 and 157 us at 2048 (0.72 of the bf16 peak).  So: one
# launch of the fused kernel while the rows fit one, the two-step route above.
TWO_STEP_MIN_ROWS = int(os.environ.get("LUTVQ_TWO_STEP_MIN_ROWS", "257"))
# Small layers (few row tiles: the fused kernel splits k and pays a fixed-order reduction of fp32 partials, while dequantize-once writes only
# 32 MB): measured on the other Llama-2-7B shapes (tools/prefill_shapes.py, profiles/r02_prefill_shapes.jsonl) fused / two-step =
# 9.6 / 18.7, 11.2 / 18.9, 23.2 / 21.1 us at 16 / 64 / 256 rows of 4096x4096 (11008x4096: 14.9 / 40.5, 16.9 / 41.2, 31.9 / 48.3; 4096x12288: 14.1 / 41.2,
# 15.2 / 41.8, 28.8 / 47.7): up to 4096x4096 weights the two-step route takes over from 129 rows.
TWO_STEP_MIN_ROWS_SMALL = int(os.environ.get("LUTVQ_TWO_STEP_MIN_ROWS_SMALL", "129"))
SMALL_LAYER_WEIGHTS = 4096 * 4096


def two_step_min_rows(out_features: int, in_features: int) -> int:
    """rows from which ops.code2x8_matmat_dequant takes the dequantize-once + library GEMM route for this layer shape"""
    return min(TWO_STEP_MIN_ROWS, TWO_STEP_MIN_ROWS_SMALL) if out_features * in_features <= SMALL_LAYER_WEIGHTS else TWO_STEP_MIN_ROWS


def code2x8_matmat_dequant(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                           bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """cuda_kernel.cpp:450-484, the prefill op: the fused tcgen05 kernel below two_step_min_rows(out, in) rows, dequantize-once +
    library GEMM from there on (whichever is faster on B200; both within the north-star tolerance of the float64 result)."""
    rows = input.numel() // max(input.size(-1), 1)
    if rows >= two_step_min_rows(codes.size(0), input.size(-1)):
        return code2x8_matmat_dequant_two_step(input, codes, codebooks, scales, bias)
    return code2x8_matmat_dequant_fused(input, codes, codebooks, scales, bias)


def code2x8_dequant(codes: torch.Tensor, codebooks: torch.Tensor, scales: Optional[torch.Tensor]) -> torch.Tensor:
    """cuda_kernel.cpp:423-448: W[out, in] = (cb0[c0] + cb1[c1]) * scales, storage dtype of the codebooks."""
    dtype = check_use_bfloat16(codebooks)
    dev = _require_cuda(codes, codebooks, scales)
    in_features = codes.size(1) * 8
    out_features = codes.size(0)
    weight = torch.empty((out_features, in_features), dtype=codebooks.dtype, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_code2x8_dequant(_ptr(codes.contiguous()), _ptr(weight), _ptr(codebooks.contiguous()),
                                          _ptr(_flat_scales(scales)), out_features, in_features, dtype, _stream(dev)))
    return weight


# ---- golden-model pair (vq.py layouts) ----
def lut_build(x: torch.Tensor, codebook: torch.Tensor, M: int, lut_mode: int = _lib.LUT_VQPY_FP16) -> torch.Tensor:
    """VQ.compute_lut (vq.py:280-285).  x fp16 [M*d]; codebook fp16 (M,N,K,d) or shared (N,K,d) -> lut (M,N,K)."""
    dev = _require_cuda(x, codebook)
    shared = codebook.dim() == 3
    N, K, d = codebook.shape[-3:]
    lut = torch.empty((M, N, K), dtype=torch.float16 if lut_mode == _lib.LUT_VQPY_FP16 else torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_lut_build(_ptr(x.contiguous()), _ptr(codebook.contiguous()), _ptr(lut), M, N, K, d,
                                    int(shared), lut_mode, _stream(dev)))
    return lut


def lut_gemv(codeword: torch.Tensor, lut: torch.Tensor, acc_mode: int = _lib.ACC_F64) -> torch.Tensor:
    """VQ.compute_lut_gemm inner loops (vq.py:293-298).  codeword uint8 (d_out,M,N); lut (M,N,K) fp16|fp32."""
    dev = _require_cuda(codeword, lut)
    d_out, M, N = codeword.shape
    K = lut.shape[-1]
    out = torch.empty(d_out, dtype=torch.float64 if acc_mode == _lib.ACC_F64 else torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_lut_gemv(_ptr(codeword.contiguous()), _ptr(lut.contiguous()), _ptr(out), d_out, M, N, K,
                                   int(lut.dtype == torch.float32), acc_mode, _stream(dev)))
    return out


def vq_dequantize(codebook: torch.Tensor, codeword: torch.Tensor, M: int) -> torch.Tensor:
    """VQ.dequantize (vq.py:269-278) -> fp16 (d_out, M*d)."""
    dev = _require_cuda(codebook, codeword)
    shared = codebook.dim() == 3
    N, K, d = codebook.shape[-3:]
    d_out = codeword.shape[0]
    W = torch.empty((d_out, M * d), dtype=torch.float16, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_vq_dequantize(_ptr(codebook.contiguous()), _ptr(codeword.contiguous()), _ptr(W), d_out, M, N,
                                        K, d, int(shared), _stream(dev)))
    return W


def code2x8_matmat_group_packed(input: torch.Tensor, packed: List[torch.Tensor], out_features: List[int], codebooks: List[torch.Tensor],
                                scales: List[Optional[torch.Tensor]], bias: Optional[List[Optional[torch.Tensor]]] = None,
                                out: Optional[List[torch.Tensor]] = None, gate: Optional[torch.Tensor] = None) -> List[torch.Tensor]:
    """n <= 3 linears that read the same input row (q/k/v, gate/up) in ONE launch (lutvq_code2x8_matmat_group_packed);
    bit-identical to n code2x8_matmat_packed calls.  With `gate` the activations are silu(gate) * input, evaluated in
    the kernel's activation load (lutvq_code2x8_matmat_group_packed_swiglu; fp16)."""
    dtype = check_use_bfloat16(input)
    n = len(packed)
    dev = _require_cuda(input, *packed, *codebooks)
    in_features = input.size(-1)
    assert input.numel() == in_features, "grouped launches are batch 1"
    x = input.reshape(-1).contiguous()
    ys = out if out is not None else [torch.empty(o, dtype=input.dtype, device=dev) for o in out_features]
    sc = [_flat_scales(s) for s in scales]
    bi = [None] * n if bias is None else [None if b is None else b.contiguous() for b in bias]
    cbs = [c.contiguous() for c in codebooks]
    arr = lambda ts: (ctypes.c_void_p * n)(*[None if t is None else t.data_ptr() for t in ts])
    pm = (ctypes.c_int * n)(*out_features)
    need = int(lib().lutvq_group_workspace_bytes(n, pm, in_features))
    key = ("group", dev.index if dev.index is not None else torch.cuda.current_device(), torch.cuda.current_stream(dev).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.zeros(max(need, 1 << 18), dtype=torch.uint8, device=dev)
        _workspaces[key] = ws
    with torch.cuda.device(dev):
        tail = (n, arr(packed), arr(cbs), arr(sc), arr(bi), arr(ys), pm, in_features, dtype, 0, _ptr(ws), ws.numel(), _stream(dev))
        if gate is None:
            check(lib().lutvq_code2x8_matmat_group_packed(_ptr(x), *tail))
        else:
            g = gate.reshape(-1).contiguous()
            assert g.numel() == in_features and g.dtype == input.dtype and g.device == x.device
            check(lib().lutvq_code2x8_matmat_group_packed_swiglu(_ptr(g), _ptr(x), *tail))
    return ys


def _kx8_fused(in_features: int, N: int) -> bool:
    """csrc/kx8.cu kx8_fused: the fused LUT-GEMV takes 1 / 2 / 4 / 8 codebooks when a code row is whole 16-byte chunks"""
    return N in (1, 2, 4, 8) and (in_features // 8 * N) % 16 == 0


def codeKx8_matmat(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                   bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """codeKx8_matmat / code1x8_matmat (cuda_kernel.cpp:521-684): any number of 256-entry codebooks, groups of 8.
    N in {1, 2, 4, 8}: the fused LUT-GEMV (fp16 / bf16); other N: the unfused fp32 LUT pair (fp16)."""
    dtype = check_use_bfloat16(input)
    dev = _require_cuda(input, codes, codebooks, scales, bias)
    N = codes.size(-1)
    in_features, out_features = input.size(-1), codes.size(0)
    flat = input.reshape(-1, in_features).contiguous()
    y = torch.empty((flat.size(0), out_features), dtype=input.dtype, device=dev)
    need = int(lib().lutvq_codeKx8_workspace_bytes(out_features, in_features, N))
    with torch.cuda.device(dev):
        if _kx8_fused(in_features, N) and flat.size(0) > 1 and BATCH_GROUPED:
            ws = batch_workspace(dev, out_features, in_features, flat.size(0))     # up to 8 rows per launch (lutvq.h)
        elif _kx8_fused(in_features, N):
            ws = workspace(dev, out_features, in_features)          # zero-initialised split-k words, handed back zero by every call
            if ws.numel() < need:
                ws = torch.zeros(need, dtype=torch.uint8, device=dev)
        else:
            ws = torch.empty(need, dtype=torch.uint8, device=dev)   # scratch LUT of the unfused path: never the shared zeroed words
        check(lib().lutvq_codeKx8_matmat(_ptr(flat), _ptr(codes.contiguous()), _ptr(codebooks.contiguous()), _ptr(_flat_scales(scales)),
                                         _ptr(None if bias is None else bias.contiguous()), _ptr(y), flat.size(0), out_features, in_features, N,
                                         dtype, _ptr(ws), ws.numel(), _stream(dev)))
    return y.reshape(tuple(input.shape[:-1]) + (out_features,))


def codeKx8_dequant(codes: torch.Tensor, codebooks: torch.Tensor, scales: Optional[torch.Tensor]) -> torch.Tensor:
    """CodeKx8Dequant (cuda_kernel.cu:392-468): W = ((cb0 + cb1) + cb2) + ... in the storage type (fp16 / bf16), then * scales."""
    dtype = check_use_bfloat16(codebooks)
    dev = _require_cuda(codes, codebooks, scales)
    O, G, N = codes.shape
    W = torch.empty((O, G * 8), dtype=codebooks.dtype, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_codeKx8_dequant(_ptr(codes.contiguous()), _ptr(W), _ptr(codebooks.contiguous()), _ptr(_flat_scales(scales)), O, G * 8, N,
                                          dtype, _stream(dev)))
    return W


def _check_1x16(codes: torch.Tensor, codebooks: torch.Tensor) -> int:
    if codes.dtype != torch.int16:
        raise TypeError(f"1x16 codes must be int16, got {codes.dtype}")
    gs = codebooks.size(3)
    if gs not in (8, 16):                                   # cuda_kernel.cpp:136-145
        raise NotImplementedError(f"AQLM CUDA kernels only support codebooks with 8 or 16 features. Got {gs}.")
    if codebooks.size(0) != 1 or codebooks.size(1) != 65536 or codebooks.size(2) != 1:
        raise NotImplementedError(f"1x16 kernels take codebooks [1, 65536, 1, 8|16], got {tuple(codebooks.shape)}")
    return gs


def code1x16_matvec(A: torch.Tensor, B: torch.Tensor, C: torch.Tensor, codebook: torch.Tensor, use_bfloat16: bool) -> None:
    """code1x16_matvec (cuda_kernel.cpp:113-146): C[prob_m] = round(W(A) . B), raw launcher semantics (no scale); A = codes.squeeze(2)."""
    dev = _require_cuda(A, B, C, codebook)
    gs = _check_1x16(A, codebook)
    with torch.cuda.device(dev):
        check(lib().lutvq_code1x16_matvec(_ptr(A.contiguous()), _ptr(B.contiguous()), _ptr(C), _ptr(codebook.contiguous()), C.size(0), B.size(0), gs,
                                          _lib.LUTVQ_BF16 if use_bfloat16 else _lib.LUTVQ_F16, _stream(dev)))


def code1x16_matmat(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                    bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """code1x16_matmat (cuda_kernel.cpp:148-179): codes int16 [O, I/gs, 1], codebooks [1, 65536, 1, gs]."""
    dtype = check_use_bfloat16(input)
    dev = _require_cuda(input, codes, codebooks, scales, bias)
    gs = _check_1x16(codes, codebooks)
    in_features, out_features = input.size(-1), codes.size(0)
    flat = input.reshape(-1, in_features).contiguous()
    y = torch.empty((flat.size(0), out_features), dtype=input.dtype, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_code1x16_matmat(_ptr(flat), _ptr(codes.contiguous()), _ptr(codebooks.contiguous()), _ptr(_flat_scales(scales)),
                                          _ptr(None if bias is None else bias.contiguous()), _ptr(y), flat.size(0), out_features, in_features, gs,
                                          dtype, _stream(dev)))
    return y.reshape(tuple(input.shape[:-1]) + (out_features,))


def code1x16_dequant(codes: torch.Tensor, codebooks: torch.Tensor, scales: Optional[torch.Tensor]) -> torch.Tensor:
    """code1x16_dequant (cuda_kernel.cpp:181-209): W fp16/bf16 [O, I] = codebook[code] * scales, bit-exact."""
    dtype = check_use_bfloat16(codebooks)
    dev = _require_cuda(codes, codebooks, scales)
    gs = _check_1x16(codes, codebooks)
    O, G = codes.size(0), codes.size(1)
    W = torch.empty((O, G * gs), dtype=codebooks.dtype, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_code1x16_dequant(_ptr(codes.contiguous()), _ptr(W), _ptr(codebooks.contiguous()), _ptr(_flat_scales(scales)), O, G * gs, gs,
                                           dtype, _stream(dev)))
    return W


def _scale_bias_unflatten(flat_output: torch.Tensor, scales: torch.Tensor, bias: Optional[torch.Tensor], input_shape) -> torch.Tensor:
    """scale_bias_unflatten_output (cuda_kernel.cpp:95-111)"""
    flat_output = flat_output * scales.flatten().unsqueeze(0)
    if bias is not None:
        flat_output = flat_output + bias.unsqueeze(0)
    return flat_output.reshape(tuple(input_shape[:-1]) + (flat_output.size(-1),)).clone()


def code1x16_matmat_dequant(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                            bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """code1x16_matmat_dequant (cuda_kernel.cpp:229-283): our dequantize kernel (no scale), then the plain library GEMM the
    reference also calls (F.linear = cuBLAS), then the reference's epilogue."""
    check_use_bfloat16(input)
    W = code1x16_dequant(codes, codebooks, None)
    flat = input.reshape(-1, input.size(-1))
    return _scale_bias_unflatten(torch.nn.functional.linear(flat, W), scales, bias, input.shape)


def code1x16_matmat_dequant_transposed(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                                       bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """code1x16_matmat_dequant_transposed (cuda_kernel.cpp:285-354): (input * scales) @ W (the backward-pass product)."""
    check_use_bfloat16(codebooks)
    W = code1x16_dequant(codes, codebooks, None)
    scaled = (input.reshape(-1, input.size(-1)) * scales.flatten().unsqueeze(0)).reshape(input.shape)
    return torch.nn.functional.linear(scaled, W.transpose(0, 1), bias)


def code2x8_matmat_dequant_transposed(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                                      bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """code2x8_matmat_dequant_transposed (cuda_kernel.cpp:486-519): (input * scales) @ W -- the input-gradient product of the
    reference's autograd op.  Our dequantize kernel (no scale) + the plain library GEMM the reference also calls."""
    check_use_bfloat16(codebooks)
    W = code2x8_dequant(codes, codebooks, None)
    scaled = (input.reshape(-1, input.size(-1)) * scales.flatten().unsqueeze(0)).reshape(input.shape)
    return torch.nn.functional.linear(scaled, W.transpose(0, 1), bias)


def code1x8_matmat(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                   bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """code1x8_matmat (cuda_kernel.cpp:550-584): one 256-entry codebook; the N = 1 case of codeKx8_matmat."""
    return codeKx8_matmat(input, codes, codebooks, scales, bias)


def code1x8_dequant(codes: torch.Tensor, codebooks: torch.Tensor, scales: Optional[torch.Tensor]) -> torch.Tensor:
    """code1x8_dequant (cuda_kernel.cpp:586-611)."""
    return codeKx8_dequant(codes, codebooks, scales)


def code1x8_matmat_dequant(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                           bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """code1x8_matmat_dequant (cuda_kernel.cpp:613-647): dequantize (our kernel) + library GEMM + the reference's epilogue."""
    check_use_bfloat16(input)
    W = codeKx8_dequant(codes, codebooks, None)
    return _scale_bias_unflatten(torch.nn.functional.linear(input.reshape(-1, input.size(-1)), W), scales, bias, input.shape)


def code1x8_matmat_dequant_transposed(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                                      bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """code1x8_matmat_dequant_transposed (cuda_kernel.cpp:649-684)."""
    check_use_bfloat16(codebooks)
    W = codeKx8_dequant(codes, codebooks, None)
    scaled = (input.reshape(-1, input.size(-1)) * scales.flatten().unsqueeze(0)).reshape(input.shape)
    return torch.nn.functional.linear(scaled, W.transpose(0, 1), bias)


def simd_lut_lookup_add(idx: torch.Tensor, lut: torch.Tensor, output: torch.Tensor, n_runs: int) -> torch.Tensor:
    """lookup_and_add_simd, _SCALAR_ semantics (benchmark/kernel/simd_lut/bench_avx_lut.cpp:26-138, :98-107):
    output[k] += lut[idx[j, k]] for every row j, repeated n_runs times, int16 wraparound.  idx uint8 [n_lut_per_run, CB],
    lut int8 [n_entries], output int16 [CB] (updated in place and returned)."""
    dev = _require_cuda(idx, lut, output)
    assert idx.dtype == torch.uint8 and lut.dtype == torch.int8 and output.dtype == torch.int16
    assert idx.dim() == 2 and idx.is_contiguous() and output.is_contiguous() and output.numel() == idx.size(1)
    cb = idx.size(1)
    need = int(lib().lutvq_simd_lut_workspace_bytes(cb))
    key = ("simd_lut", dev.index if dev.index is not None else torch.cuda.current_device())
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.zeros(max(need, 1024), dtype=torch.uint8, device=dev)
        _workspaces[key] = ws
    with torch.cuda.device(dev):
        check(lib().lutvq_simd_lut_lookup_add(_ptr(idx), _ptr(lut.contiguous()), _ptr(output), cb, lut.numel(), int(n_runs), idx.size(0),
                                              _ptr(ws), ws.numel(), _stream(dev)))
    return output


# ---- T-MAC bit-serial integer LUT (benchmark/kernel/bs_lut_gemm/tbl.cc, lut_ctor.cc; include/lutvq.h) ----
def _tmac_check(t: torch.Tensor, dtype: torch.dtype, name: str) -> torch.Tensor:
    if t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.contiguous()


def tbl_g4_int8_int32_update(m: int, c: torch.Tensor, lut: torch.Tensor, a: torch.Tensor, K: int) -> torch.Tensor:
    """tbl_g4_int8_int32_update_impl<K, Bits> (tbl.cc:536-630): c int32[m] accumulated in place and returned."""
    dev = _require_cuda(c, lut, a)
    assert c.is_contiguous() and c.dtype == torch.int32 and c.numel() >= m
    with torch.cuda.device(dev):
        check(lib().lutvq_tbl_g4_int8_int32_update(m, _ptr(c), _ptr(_tmac_check(lut, torch.int8, "lut")), _ptr(_tmac_check(a, torch.uint8, "a")),
                                                   K, _stream(dev)))
    return c


def tbl_g4_int8_float_update(m: int, c: torch.Tensor, lut: torch.Tensor, a: torch.Tensor, scales: torch.Tensor, lut_scales: torch.Tensor,
                             lut_biases: torch.Tensor, K: int, bits: int, act_k: int, fast_aggregation: bool = False,
                             zero_point: bool = False) -> torch.Tensor:
    """tbl_g4_int8_float_update_impl<has_scale, K, Bits, ActK, FastAggregation, ZeroPoint, OneScale> (tbl.cc:323-532), x86 path."""
    dev = _require_cuda(c, lut, a, scales, lut_scales, lut_biases)
    assert c.is_contiguous() and c.dtype == torch.float32 and c.numel() >= m
    f32 = torch.float32
    with torch.cuda.device(dev):
        check(lib().lutvq_tbl_g4_int8_float_update(m, _ptr(c), _ptr(_tmac_check(lut, torch.int8, "lut")), _ptr(_tmac_check(a, torch.uint8, "a")),
                                                   _ptr(_tmac_check(scales, f32, "scales")), _ptr(_tmac_check(lut_scales, f32, "lut_scales")),
                                                   _ptr(_tmac_check(lut_biases, f32, "lut_biases")), K, bits, act_k, int(fast_aggregation),
                                                   int(zero_point), _stream(dev)))
    return c


def tbl_reset(c: torch.Tensor) -> torch.Tensor:
    """tbl_{int8,int16,int32,float}_reset (tbl.cc:697-715)."""
    dev = _require_cuda(c)
    fn = {torch.int8: "lutvq_tbl_int8_reset", torch.int16: "lutvq_tbl_int16_reset", torch.int32: "lutvq_tbl_int32_reset",
          torch.float32: "lutvq_tbl_float_reset"}[c.dtype]
    with torch.cuda.device(dev):
        check(getattr(lib(), fn)(c.numel(), _ptr(c), _stream(dev)))
    return c


def partial_max_reset(lut_scales: torch.Tensor) -> torch.Tensor:
    dev = _require_cuda(lut_scales)
    with torch.cuda.device(dev):
        check(lib().lutvq_partial_max_reset(_ptr(lut_scales), _stream(dev)))
    return lut_scales


def partial_max_g4_int8_k8(lut_scales: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """partial_max_g4_int8_k8 (lut_ctor.cc:232-259) over the 32 activations at b; lut_scales: float32 view of ONE element."""
    dev = _require_cuda(lut_scales, b)
    assert lut_scales.dtype == torch.float32 and b.dtype == torch.float32 and b.is_contiguous() and b.numel() >= 32
    with torch.cuda.device(dev):
        check(lib().lutvq_partial_max_g4_int8_k8(_ptr(lut_scales), _ptr(b), _stream(dev)))
    return lut_scales


def lut_ctor_g4_int8(act_k: int, b: torch.Tensor, lut_scales: torch.Tensor):
    """lut_ctor_g4_int8_impl (lut_ctor.cc:38-221) -> (qlut int8 [act_k/4, 16], lut_scales, lut_biases float32 [1])."""
    dev = _require_cuda(b, lut_scales)
    assert b.dtype == torch.float32 and b.is_contiguous() and b.numel() >= act_k and lut_scales.dtype == torch.float32
    qlut = torch.empty((act_k // 4, 16), dtype=torch.int8, device=dev)
    biases = torch.zeros(1, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_lut_ctor_g4_int8(act_k, _ptr(qlut), _ptr(b), _ptr(lut_scales), _ptr(biases), _stream(dev)))
    return qlut, lut_scales, biases


def tmac_preprocessor_int8(B: torch.Tensor):
    """preprocessor_t1_int8_* (T-MAC deploy/tuned/kernels.cc:1002-1037) -> (qlut int8 [K/4, 16], lut_scales [K/32], lut_biases [K/32])."""
    dev = _require_cuda(B)
    assert B.dtype == torch.float32 and B.dim() == 1 and B.is_contiguous()
    K = B.numel()
    qlut = torch.empty((K // 4, 16), dtype=torch.int8, device=dev)
    ls = torch.empty(K // 32, dtype=torch.float32, device=dev)
    lb = torch.empty(K // 32, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_tmac_preprocessor_int8(K, _ptr(B), _ptr(ls), _ptr(lb), _ptr(qlut), _stream(dev)))
    return qlut, ls, lb


def tmac_qgemm_lut_int8(M: int, K: int, bits: int, bm: int, A: torch.Tensor, qlut: torch.Tensor, scales: torch.Tensor,
                        lut_scales: torch.Tensor, lut_biases: torch.Tensor) -> torch.Tensor:
    """qgemm_lut_t1_int8_m{bm}_k{K}_n1_b{bits} (T-MAC deploy/tuned/kernels.cc:827-960) -> C float32 [M]."""
    dev = _require_cuda(A, qlut, scales, lut_scales, lut_biases)
    f32 = torch.float32
    C = torch.empty(M, dtype=f32, device=dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_tmac_qgemm_lut_int8(M, K, bits, bm, _ptr(_tmac_check(A, torch.uint8, "A")), _ptr(_tmac_check(qlut, torch.int8, "qlut")),
                                              _ptr(_tmac_check(scales, f32, "scales")), _ptr(_tmac_check(lut_scales, f32, "lut_scales")),
                                              _ptr(_tmac_check(lut_biases, f32, "lut_biases")), _ptr(C), _stream(dev)))
    return C
