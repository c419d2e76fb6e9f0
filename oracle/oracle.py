"""CPU oracle for the LUT-VQ hot path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product path
(``rvv-lut-vq_b200/`` + ``liblutvq.so``) never touches ``oracle/``.

Two layers, both restating the reference (paths relative to /root/reference):

* numpy restatements (``np_*``) that follow the reference line by line and are small-case
  checkers: ``vq_dataflow_sim/vq.py:269-307`` and
  ``benchmark/AQLM/inference_lib/src/aqlm/{utils.py:24-70, inference_kernels/numba_kernel.py:38-48}``;
* ctypes bindings (``c_*``) to ``liblutvq_oracle.so`` (``lutvq_oracle.c``), the fast C restatement
  used at full sizes and as the timed CPU baseline.

Parity: PINNED -- ``oracle/gen_golden.py`` ran the reference's own Python in the build container
and froze its outputs under ``tests/golden/``; ``tests/test_oracle_golden.py`` replays them.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liblutvq_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the C oracle (and, when /root/reference exists, oracle/_ref). Idempotent."""
    if force or not os.path.exists(_LIB_PATH) or any(
        os.path.exists(os.path.join(_HERE, s)) and os.path.getmtime(os.path.join(_HERE, s)) > os.path.getmtime(_LIB_PATH)
        for s in ("lutvq_oracle.c", "tmac_oracle.c", "pq4_oracle.c")
    ):
        subprocess.check_call(["make", "-C", _HERE, "-B", "all"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference") and os.path.exists(os.path.join(_HERE, "ref_wrap_tmac.cc")):
        subprocess.call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.orc_num_threads.restype = ctypes.c_int
        P, I, U64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint64
        sigs = {
            "orc_lut_build_vqpy": [P, P, I, I, I, I, I, P],
            "orc_lut_build_f32": [P, P, I, I, I, I, P],
            "orc_lut_gemv_f64": [P, P, I, I, I, I, P],
            "orc_dequantize": [P, I, P, P, I, I, I, I, I, I, P],
            "orc_dense_gemv_f64": [P, P, I, I, I, P],
            "orc_dense_gemm_f64": [P, P, I, I, I, I, P],
            "orc_aqlm_gemv_f64": [P, P, P, I, I, I, I, I, P],
            "orc_aqlm_gemv_f32": [P, P, P, P, P, I, I, I, I, P, P],
            "orc_aqlm_gemv_rowmajor": [P, P, P, P, P, I, I, I, I, I, P, P],
            "orc_scale_bias_epilogue": [P, P, P, I, I, I],
            "orc_unpack_codes": [P, ctypes.c_size_t, P],
            "orc_simd_lut_scalar": [P, P, P, I, U64, U64],
            "orc_set_num_threads": [I],
            "orc_tbl_g4_int8_int32_update": [I, P, P, P, I],
            "orc_tbl_g4_int8_float_update": [I, P, P, P, P, P, P, I, I, I, I, I],
            "orc_partial_max_g4_int8_k8": [P, P],
            "orc_lut_ctor_g4_int8": [I, P, P, P, P],
            "orc_tmac_preprocessor_int8": [I, P, P, P, P],
            "orc_tmac_qgemm_lut_int8": [I, I, I, I, P, P, P, P, P, P],
        }
        for name, args in sigs.items():
            fn = getattr(_lib, name)
            fn.argtypes, fn.restype = args, None
        Z = ctypes.c_size_t
        for name, args in {
            "orc_pq4_pack_codes": [P, Z, Z, Z, Z, Z, P], "orc_pq4_pack_codes_range": [P, Z, Z, Z, Z, Z, P],
            "orc_pq4_get_packed_element": [P, Z, Z, Z, Z], "orc_pq4_set_packed_element": [P, I, Z, Z, Z, Z],
            "orc_pq4_pack_LUT": [I, I, P, P], "orc_pq4_pack_LUT_qbs": [I, I, P, P], "orc_pq4_pack_LUT_qbs_q_map": [I, I, P, P, P],
            "orc_pq4_accumulate_loop": [I, Z, I, I, P, P, P, I],
        }.items():
            fn = getattr(_lib, name)
            fn.argtypes, fn.restype = args, ctypes.c_int
    return _lib


def _p(a: Optional[np.ndarray]):
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"], "oracle expects C-contiguous arrays"
    return a.ctypes.data_as(ctypes.c_void_p)


def num_threads() -> int:
    return int(lib().orc_num_threads())


def set_num_threads(n: int) -> None:
    lib().orc_set_num_threads(int(n))


def calibrate_threads(fn, max_threads: int = 0) -> int:
    """Pick the OpenMP thread count (1, 2, 4, ... host cores) at which `fn()` runs fastest and leave it set.
    (On oversubscribed VMs more threads can be slower; the baseline should be the best the host can do.)"""
    import time
    hi = max_threads or (os.cpu_count() or 1)
    cands, t = [], 1
    while t < hi:
        cands.append(t)
        t *= 2
    cands.append(hi)
    best, best_t = 1, float("inf")
    for n in cands:
        set_num_threads(n)
        fn()
        t0 = time.perf_counter()
        fn(); fn()
        dt = (time.perf_counter() - t0) / 2
        if dt < best_t:
            best, best_t = n, dt
    set_num_threads(best)
    return best


# --------------------------------------------------------------------------------------------
# helpers: dtype views.  fp16 lives in numpy natively; bf16 is carried as uint16 bit patterns.
# --------------------------------------------------------------------------------------------
def f32_to_bf16_bits(a: np.ndarray) -> np.ndarray:
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32).astype(np.uint64)
    r = (u + 0x7FFF + ((u >> 16) & 1)) >> 16
    return r.astype(np.uint16)


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    return (b.astype(np.uint32) << 16).view(np.float32)


def to_bits(a: np.ndarray, dtype: int) -> np.ndarray:
    """storage bits (uint16) of an array given as float: dtype 0 = fp16, 1 = bf16"""
    if a.dtype == np.uint16:
        return np.ascontiguousarray(a)
    if dtype == 0:
        return np.ascontiguousarray(a.astype(np.float16)).view(np.uint16)
    return f32_to_bf16_bits(a.astype(np.float32))


def from_bits(b: np.ndarray, dtype: int) -> np.ndarray:
    return b.view(np.float16).astype(np.float32) if dtype == 0 else bf16_bits_to_f32(b)


# --------------------------------------------------------------------------------------------
# numpy restatements (small cases; line-by-line with the reference)
# --------------------------------------------------------------------------------------------
def np_unpack_int_data(codes: np.ndarray, nbits: int = 8) -> np.ndarray:
    """utils.py:29-31: data.to(int64) % 2**nbits"""
    return codes.astype(np.int64) % (2**nbits)


def np_pack_int_data(idx: np.ndarray, nbits: int = 8) -> np.ndarray:
    """utils.py:24-26 (without mutating the input): values >= 2^(nbits-1) wrap negative, stored in int8"""
    d = idx.astype(np.int64).copy()
    d[d >= 2 ** (nbits - 1)] -= 2**nbits
    return d.astype(np.int8 if nbits <= 8 else np.int16)


def np_compute_lut(x: np.ndarray, codebook: np.ndarray) -> np.ndarray:
    """vq.py:280-285.  x (D,) ; codebook (M,N,K,d) fp16 -> lut (M,N,K) fp16 (numpy fp16 semantics)."""
    M, N, K, d = codebook.shape
    xr = x.reshape(M, 1, 1, d).astype(np.float16)
    return (codebook * xr).sum(axis=-1)


def np_compute_lut_gemm(x: np.ndarray, codebook: np.ndarray, codeword: np.ndarray) -> np.ndarray:
    """vq.py:287-302 vectorised: out[i,j] = sum_{m,n} lut_i[m,n,code[j,m,n]] in float64."""
    M, N, K, d = codebook.shape
    out = np.zeros((x.shape[0], codeword.shape[0]))
    mi = np.arange(M)[None, :, None]
    ni = np.arange(N)[None, None, :]
    for i in range(x.shape[0]):
        lut = np_compute_lut(x[i], codebook)
        out[i] = lut[mi, ni, codeword].astype(np.float64).sum(axis=(1, 2))
    return out


def np_dequantize(codebook: np.ndarray, codeword: np.ndarray) -> np.ndarray:
    """vq.py:269-278: weight[j,m,:] = sum_n codebook[m,n,code[j,m,n],:] (float64) -> fp16 (d_out, D)."""
    M, N, K, d = codebook.shape
    d_out = codeword.shape[0]
    w = np.zeros((d_out, M, d))
    mi = np.arange(M)[None, :]
    for n in range(N):
        w += codebook[mi, n, codeword[:, :, n], :]
    return w.reshape(d_out, M * d).astype(np.float16)


def np_aqlm_dequantize_weight(codes_idx: np.ndarray, codebooks: np.ndarray, scales: Optional[np.ndarray]) -> np.ndarray:
    """utils.py:43-70 for out_group_size=1: W[o, g*d:(g+1)*d] = (sum_n codebooks[n, idx[o,g,n], 0, :]) * scales[o].
    Computed in float32 from the given arrays (embedding_bag accumulates in the storage dtype; callers
    compare by tolerance except for N=2 where a single add is exact in both)."""
    O, G, N = codes_idx.shape
    d = codebooks.shape[-1]
    w = np.zeros((O, G, d), dtype=np.float32)
    for n in range(N):
        w += codebooks[n, codes_idx[:, :, n], 0, :].astype(np.float32)
    if scales is not None:
        w = w * scales.reshape(O, 1, 1).astype(np.float32)
    return w.reshape(O, G * d)


def np_aqlm_1x16_dequantize(codes_i16: np.ndarray, codebook_f16: np.ndarray, scales_f16: Optional[np.ndarray]) -> np.ndarray:
    """AQLM 1x16 through utils.py:29-31 (unpack) and :43-70 (_dequantize_weight, one codebook: a pure gather, then
    `* scales` = one fp16 RN multiply) = Code1x16Dequant (cuda_kernel.cu:98-142) + `weight *= scales` (cuda_kernel.cpp:206).
    codes int16 [O, G, 1] or [O, G]; codebook fp16 [1, 65536, 1, gs]; -> fp16 [O, G * gs], bit-exact."""
    idx = np_unpack_int_data(codes_i16.reshape(codes_i16.shape[0], -1), 16)
    cb = codebook_f16.reshape(65536, -1)
    w = cb[idx]                                                    # [O, G, gs]
    if scales_f16 is not None:
        w = (w.astype(np.float32) * scales_f16.reshape(-1, 1, 1).astype(np.float32)).astype(np.float16)   # exact product, one rounding
    return np.ascontiguousarray(w.reshape(w.shape[0], -1))


def codebook_1x16(seed: int, gs: int) -> np.ndarray:
    """the 65536-entry codebook of a 1x16 fixture, regenerated from its seed (same code as oracle/gen_golden.py)"""
    return np.random.default_rng(seed).standard_normal((1, 65536, 1, gs)).astype(np.float16)


def np_numba_gemv_lut(x: np.ndarray, codebooks: np.ndarray, codes_alt: np.ndarray, scales: np.ndarray) -> np.ndarray:
    """numba_kernel.py:38-48 restated with numpy (fp32).  codes_alt uint8 [G, O, N]."""
    N, K, _, d = codebooks.shape
    G, O, _ = codes_alt.shape
    lut = x.reshape(-1, d).astype(np.float32) @ codebooks.reshape(-1, d).astype(np.float32).T
    lut = lut.reshape(G, N, K)
    gi = np.arange(G)[:, None, None]
    ci = np.arange(N)[None, None, :]
    out = lut[gi, ci, codes_alt].sum(axis=(0, 2), dtype=np.float32)
    return out * scales.reshape(-1).astype(np.float32)


# --------------------------------------------------------------------------------------------
# ctypes bindings to the C restatement
# --------------------------------------------------------------------------------------------
def c_lut_build_vqpy(x_f16: np.ndarray, codebook_f16: np.ndarray, M: int) -> np.ndarray:
    """codebook (M,N,K,d) or shared (N,K,d).  returns lut fp16 (M,N,K)"""
    shared = codebook_f16.ndim == 3
    N, K, d = codebook_f16.shape[-3:]
    x = np.ascontiguousarray(x_f16, dtype=np.float16)
    cb = np.ascontiguousarray(codebook_f16, dtype=np.float16)
    lut = np.empty((M, N, K), dtype=np.float16)
    lib().orc_lut_build_vqpy(_p(x), _p(cb), int(shared), M, N, K, d, _p(lut))
    return lut


def c_lut_gemv_f64(lut_f16: np.ndarray, codeword_u8: np.ndarray) -> np.ndarray:
    M, N, K = lut_f16.shape
    d_out = codeword_u8.shape[0]
    out = np.empty(d_out, dtype=np.float64)
    lib().orc_lut_gemv_f64(_p(np.ascontiguousarray(lut_f16)), _p(np.ascontiguousarray(codeword_u8)), d_out, M, N, K, _p(out))
    return out


def c_dequantize(codebook_bits: np.ndarray, codeword_u8: np.ndarray, scales_bits: Optional[np.ndarray], M: int, dtype: int = 0) -> np.ndarray:
    """returns W storage bits uint16 (d_out, M*d).  codebook (M,N,K,d) or shared (N,K,d), uint16 bits or fp16."""
    cb = codebook_bits.view(np.uint16) if codebook_bits.dtype != np.uint16 else codebook_bits
    cb = np.ascontiguousarray(cb)
    shared = cb.ndim == 3
    N, K, d = cb.shape[-3:]
    cw = np.ascontiguousarray(codeword_u8, dtype=np.uint8)
    d_out = cw.shape[0]
    sc = None
    if scales_bits is not None:
        sc = np.ascontiguousarray(scales_bits.view(np.uint16) if scales_bits.dtype != np.uint16 else scales_bits).reshape(-1)
    W = np.empty((d_out, M * d), dtype=np.uint16)
    lib().orc_dequantize(_p(cb), int(shared), _p(cw), _p(sc), d_out, M, N, K, d, dtype, _p(W))
    return W


def c_dense_gemv_f64(x_bits: np.ndarray, W_bits: np.ndarray, dtype: int = 0) -> np.ndarray:
    d_out, D = W_bits.shape
    out = np.empty(d_out, dtype=np.float64)
    lib().orc_dense_gemv_f64(_p(np.ascontiguousarray(x_bits.view(np.uint16))), _p(np.ascontiguousarray(W_bits.view(np.uint16))), d_out, D, dtype, _p(out))
    return out


def c_dense_gemm_f64(x_bits: np.ndarray, W_bits: np.ndarray, dtype: int = 0) -> np.ndarray:
    B, I = x_bits.shape
    O = W_bits.shape[0]
    out = np.empty((B, O), dtype=np.float64)
    lib().orc_dense_gemm_f64(_p(np.ascontiguousarray(x_bits.view(np.uint16))), _p(np.ascontiguousarray(W_bits.view(np.uint16))), B, O, I, dtype, _p(out))
    return out


def c_aqlm_gemv_f64(x_bits: np.ndarray, codes_i8: np.ndarray, codebooks_bits: np.ndarray, dtype: int = 0) -> np.ndarray:
    """exact y64 from raw operands.  codes int8 [O,G,N]; codebooks [N,256,1,d] or [N,256,d] storage bits."""
    O, G, N = codes_i8.shape
    d = codebooks_bits.shape[-1]
    out = np.empty(O, dtype=np.float64)
    lib().orc_aqlm_gemv_f64(
        _p(np.ascontiguousarray(x_bits.view(np.uint16)).reshape(-1)),
        _p(np.ascontiguousarray(codes_i8.view(np.int8))),
        _p(np.ascontiguousarray(codebooks_bits.view(np.uint16))),
        O, G, N, d, dtype, _p(out))
    return out


def c_aqlm_gemv_f32(x_f32: np.ndarray, codes_alt_u8: np.ndarray, codebooks_f32: np.ndarray, scales_f32: np.ndarray,
                    bias_f32: Optional[np.ndarray] = None) -> np.ndarray:
    """numba_gemv_lut_ port.  codes_alt uint8 [G,O,N]"""
    G, O, N = codes_alt_u8.shape
    d = codebooks_f32.shape[-1]
    lut = np.empty((G, N, 256), dtype=np.float32)
    out = np.empty(O, dtype=np.float32)
    lib().orc_aqlm_gemv_f32(
        _p(np.ascontiguousarray(x_f32, dtype=np.float32).reshape(-1)),
        _p(np.ascontiguousarray(codes_alt_u8.view(np.uint8))),
        _p(np.ascontiguousarray(codebooks_f32, dtype=np.float32)),
        _p(np.ascontiguousarray(scales_f32, dtype=np.float32).reshape(-1)),
        _p(None if bias_f32 is None else np.ascontiguousarray(bias_f32, dtype=np.float32)),
        O, G, N, d, _p(lut), _p(out))
    return out


class RowMajorGemv:
    """The timed CPU baseline: preallocates scratch so repeated calls time only the math."""

    def __init__(self, O: int, G: int, N: int = 2, d: int = 8, dtype: int = 0):
        self.O, self.G, self.N, self.d, self.dtype = O, G, N, d, dtype
        self.lut = np.empty((G, N, 256), dtype=np.float32)
        self.y = np.empty(O, dtype=np.uint16)
        self._fn = lib().orc_aqlm_gemv_rowmajor

    def __call__(self, x_bits, codes_i8, codebooks_bits, scales_bits=None, bias_bits=None) -> np.ndarray:
        self._fn(_p(x_bits), _p(codes_i8), _p(codebooks_bits), _p(scales_bits), _p(bias_bits),
                 self.O, self.G, self.N, self.d, self.dtype, _p(self.lut), _p(self.y))
        return self.y


def c_aqlm_gemv_rowmajor(x_bits, codes_i8, codebooks_bits, scales_bits=None, bias_bits=None, dtype: int = 0) -> np.ndarray:
    O, G, N = codes_i8.shape
    d = codebooks_bits.shape[-1]
    f = RowMajorGemv(O, G, N, d, dtype)
    u16 = lambda a: None if a is None else np.ascontiguousarray(a.view(np.uint16)).reshape(-1)
    return f(u16(x_bits), np.ascontiguousarray(codes_i8.view(np.int8)), u16(codebooks_bits), u16(scales_bits), u16(bias_bits)).copy()


def c_scale_bias_epilogue(y_bits: np.ndarray, scales_bits, bias_bits, dtype: int = 0) -> np.ndarray:
    y = np.ascontiguousarray(y_bits.view(np.uint16)).copy()
    B, O = y.reshape(-1, y.shape[-1]).shape
    u16 = lambda a: None if a is None else np.ascontiguousarray(a.view(np.uint16)).reshape(-1)
    lib().orc_scale_bias_epilogue(_p(y), _p(u16(scales_bits)), _p(u16(bias_bits)), B, O, dtype)
    return y


def c_simd_lut_scalar(idx_u8: np.ndarray, lut_i8: np.ndarray, n_runs: int, out0: Optional[np.ndarray] = None) -> np.ndarray:
    """bench_avx_lut.cpp:98-107 scalar semantics.  idx [n_lut_per_run, CB] uint8; returns int16[CB] (accumulated onto out0)."""
    n_lut, cb = idx_u8.shape
    out = np.zeros(cb, dtype=np.int16) if out0 is None else np.ascontiguousarray(out0, dtype=np.int16).copy()
    lib().orc_simd_lut_scalar(_p(np.ascontiguousarray(idx_u8)), _p(np.ascontiguousarray(lut_i8)), _p(out), cb, n_runs, n_lut)
    return out


# --------------------------------------------------------------------------------------------
# T-MAC bit-serial integer-LUT kernels (tmac_oracle.c; reference benchmark/kernel/bs_lut_gemm/{tbl,lut_ctor}.cc)
# --------------------------------------------------------------------------------------------
def c_tbl_g4_int8_int32_update(m: int, c_i32: np.ndarray, lut_i8: np.ndarray, a_u8: np.ndarray, K: int) -> np.ndarray:
    """tbl.cc:586-627; returns the updated copy of c"""
    c = np.ascontiguousarray(c_i32, dtype=np.int32).copy()
    lib().orc_tbl_g4_int8_int32_update(m, _p(c), _p(np.ascontiguousarray(lut_i8, dtype=np.int8)), _p(np.ascontiguousarray(a_u8, dtype=np.uint8)), K)
    return c


def c_tbl_g4_int8_float_update(m, c_f32, lut_i8, a_u8, scales, lut_scales, lut_biases, K, bits, act_k, fast_aggregation=False, zero_point=False):
    """tbl.cc:433-520; returns the updated copy of c"""
    c = np.ascontiguousarray(c_f32, dtype=np.float32).copy()
    f = lambda v: np.ascontiguousarray(v, dtype=np.float32)
    lib().orc_tbl_g4_int8_float_update(m, _p(c), _p(np.ascontiguousarray(lut_i8, dtype=np.int8)), _p(np.ascontiguousarray(a_u8, dtype=np.uint8)),
                                       _p(f(scales)), _p(f(lut_scales)), _p(f(lut_biases)), K, bits, act_k, int(fast_aggregation), int(zero_point))
    return c


def c_partial_max_g4_int8_k8(lut_scale: float, b_f32: np.ndarray) -> float:
    """lut_ctor.cc:232-259"""
    s = np.array([lut_scale], dtype=np.float32)
    lib().orc_partial_max_g4_int8_k8(_p(s), _p(np.ascontiguousarray(b_f32, dtype=np.float32)))
    return float(s[0])


def c_lut_ctor_g4_int8(act_k: int, b_f32: np.ndarray, lut_scale: float):
    """lut_ctor.cc:120-216 -> (qlut int8 [act_k/32 * 8][16], lut_scales, lut_biases)"""
    qlut = np.zeros((act_k // 32 * 8, 16), dtype=np.int8)
    s, bi = np.array([lut_scale], dtype=np.float32), np.zeros(1, dtype=np.float32)
    lib().orc_lut_ctor_g4_int8(act_k, _p(qlut), _p(np.ascontiguousarray(b_f32, dtype=np.float32)), _p(s), _p(bi))
    return qlut, float(s[0]), float(bi[0])


def c_tmac_preprocessor_int8(B_f32: np.ndarray):
    """T-MAC deploy/tuned/kernels.cc:1002-1037 -> (qlut int8 [K/4][16], lut_scales [K/32], lut_biases [K/32])"""
    K = B_f32.shape[0]
    qlut = np.zeros((K // 4, 16), dtype=np.int8)
    ls, lb = np.zeros(K // 32, dtype=np.float32), np.zeros(K // 32, dtype=np.float32)
    lib().orc_tmac_preprocessor_int8(K, _p(np.ascontiguousarray(B_f32, dtype=np.float32)), _p(ls), _p(lb), _p(qlut))
    return qlut, ls, lb


def c_tmac_qgemm_lut_int8(M: int, K: int, bits: int, bm: int, A_u8, qlut_i8, scales, lut_scales, lut_biases) -> np.ndarray:
    """T-MAC deploy/tuned/kernels.cc:827-960 (float_type = float), all m tiles"""
    C = np.zeros(M, dtype=np.float32)
    f = lambda v: np.ascontiguousarray(v, dtype=np.float32)
    lib().orc_tmac_qgemm_lut_int8(M, K, bits, bm, _p(np.ascontiguousarray(A_u8, dtype=np.uint8)), _p(np.ascontiguousarray(qlut_i8, dtype=np.int8)),
                                  _p(f(scales)), _p(f(lut_scales)), _p(f(lut_biases)), _p(C))
    return C


# --------------------------------------------------------------------------------------------
# PQ4 fast-scan (pq4_oracle.c; reference benchmark/kernel/faiss-simd/pq4_fast_scan{,_search_1}.cpp)
# --------------------------------------------------------------------------------------------
def aligned_u8(n: int, align: int = 32) -> np.ndarray:
    """zeroed uint8[n] whose data pointer is `align`-byte aligned (the reference checks is_aligned_pointer)"""
    raw = np.zeros(n + align, dtype=np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n]


def pq4_nsq(M: int) -> int:
    return (M + 1) // 2 * 2


def c_pq4_pack_codes(codes_u8: np.ndarray, M: int, nb: int, bbs: int) -> np.ndarray:
    """pq4_fast_scan.cpp:45-91.  codes [ntotal, (M+1)//2] uint8 (two 4-bit codes per byte) -> blocks uint8[nb * nsq / 2], aligned"""
    codes_u8 = np.ascontiguousarray(codes_u8, dtype=np.uint8)
    nsq = pq4_nsq(M)
    blocks = aligned_u8(nb * nsq // 2)
    if lib().orc_pq4_pack_codes(_p(codes_u8), codes_u8.shape[0], M, nb, bbs, nsq, _p(blocks)):
        raise ValueError("pq4_pack_codes: bbs % 32, nb % bbs or nsq % 2 (pq4_fast_scan.cpp:53-55)")
    return blocks


def c_pq4_pack_codes_range(codes_u8: np.ndarray, M: int, i0: int, i1: int, bbs: int, blocks: np.ndarray) -> np.ndarray:
    """pq4_fast_scan.cpp:93-137: ORs vectors [i0, i1) into `blocks` (modified in place and returned)"""
    lib().orc_pq4_pack_codes_range(_p(np.ascontiguousarray(codes_u8, dtype=np.uint8)), M, i0, i1, bbs, pq4_nsq(M), _p(blocks))
    return blocks


def c_pq4_get_packed_element(blocks: np.ndarray, bbs: int, nsq: int, vector_id: int, sq: int) -> int:
    return lib().orc_pq4_get_packed_element(_p(blocks), bbs, nsq, vector_id, sq)


def c_pq4_set_packed_element(blocks: np.ndarray, code: int, bbs: int, nsq: int, vector_id: int, sq: int) -> None:
    lib().orc_pq4_set_packed_element(_p(blocks), code, bbs, nsq, vector_id, sq)


def c_pq4_pack_LUT(lut_u8: np.ndarray) -> np.ndarray:
    """pq4_fast_scan.cpp:253-264.  lut [nq, nsq, 16] uint8 -> packed uint8[nq * nsq * 16], aligned"""
    nq, nsq, _ = lut_u8.shape
    dest = aligned_u8(nq * nsq * 16)
    lib().orc_pq4_pack_LUT(nq, nsq, _p(np.ascontiguousarray(lut_u8, dtype=np.uint8)), _p(dest))
    return dest


def c_pq4_pack_LUT_qbs(qbs: int, lut_u8: np.ndarray, q_map: Optional[np.ndarray] = None) -> np.ndarray:
    """pq4_fast_scan.cpp:266-281 / 305-322"""
    nq, nsq, _ = lut_u8.shape
    n_out = sum((qbs >> s) & 15 for s in range(0, 32, 4))
    dest = aligned_u8(n_out * nsq * 16)
    src = np.ascontiguousarray(lut_u8, dtype=np.uint8)
    if q_map is None:
        got = lib().orc_pq4_pack_LUT_qbs(qbs, nsq, _p(src), _p(dest))
    else:
        got = lib().orc_pq4_pack_LUT_qbs_q_map(qbs, nsq, _p(src), _p(np.ascontiguousarray(q_map, dtype=np.int32)), _p(dest))
    if got < 0:
        raise ValueError("pq4_pack_LUT_qbs: nsq % 2 (pq4_fast_scan.cpp:267)")
    assert got == n_out
    return dest


def c_pq4_accumulate_loop(nq: int, nb: int, bbs: int, nsq: int, blocks: np.ndarray, lut_packed: np.ndarray, scale: int = -1) -> np.ndarray:
    """pq4_fast_scan_search_1.cpp:137-222 with a store-to-memory handler -> uint16 [nq, nb]; scale < 0 = no NormTableScaler"""
    accu = np.zeros((nq, nb), dtype=np.uint16)
    if lib().orc_pq4_accumulate_loop(nq, nb, bbs, nsq, _p(blocks), _p(lut_packed), _p(accu), scale):
        raise ValueError(f"pq4_accumulate_loop: alignment / bbs / nb, or nq={nq} bbs={bbs} not instantiated (pq4_fast_scan_search_1.cpp:147-168)")
    return accu


def np_pq4_distances(codes_u8: np.ndarray, lut_u8: np.ndarray, M: int, scale: int = -1) -> np.ndarray:
    """what the packed pipeline must equal, from the UNPACKED codes: dis[q, j] = sum_sq lut[q, sq, code[j, sq]] mod 2**16,
    the last two sub-quantizers times `scale` when a scaler is given"""
    nq, nsq, _ = lut_u8.shape
    n = codes_u8.shape[0]
    nib = np.zeros((n, nsq), dtype=np.int64)
    nib[:, 0:2 * codes_u8.shape[1]:2] = codes_u8 & 15
    nib[:, 1:2 * codes_u8.shape[1]:2] = codes_u8 >> 4
    if M < nsq:
        nib[:, M:] = nib[:, M:]          # an odd M keeps whatever the high nibble of the last byte holds (the reference does too)
    w = np.ones(nsq, dtype=np.int64)
    if scale >= 0:
        w[nsq - 2:] = scale
    vals = lut_u8.astype(np.int64)[:, np.arange(nsq)[None, :], nib]          # [nq, n, nsq]
    return ((vals * w).sum(-1) & 0xFFFF).astype(np.uint16)


# ---- the reference's own compiled kernels (oracle/_ref, built by `make -C oracle ref` where /root/reference exists) ----
def ref_tmac():
    path = os.path.join(_HERE, "_ref", "libtmac_ref.so")
    return ctypes.CDLL(path) if os.path.exists(path) else None


def ref_pq4():
    path = os.path.join(_HERE, "_ref", "libpq4_ref.so")
    return ctypes.CDLL(path) if os.path.exists(path) else None


def ref_simd_lut():
    path = os.path.join(_HERE, "_ref", "libsimdlut_ref.so")
    return ctypes.CDLL(path) if os.path.exists(path) else None
