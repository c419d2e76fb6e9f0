"""CPU: the oracle (oracle/lutvq_oracle.c + oracle/oracle.py) replayed against the golden vectors that
oracle/gen_golden.py froze from the reference's own Python (vq.py, aqlm utils, numba_gemm_lut).
This is what pins the oracle (the reference ships no vectors of its own)."""
import os

import numpy as np
import pytest

VQ_CASES = ["vq_d128_m16_n2_k256", "vq_d64_m8_n1_k256", "vq_d64_m8_n4_k256", "vq_d96_m6_n2_k256_d16"]
AQLM_CASES = ["aqlm_2x8_i256_o96", "aqlm_2x8_i1024_o64", "aqlm_1x8_i128_o32", "aqlm_4x8_i128_o48"]


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


@pytest.mark.parametrize("name", VQ_CASES)
def test_vqpy_lut_build_bit_exact(orc, golden_dir, name):
    g = _load(golden_dir, name)
    M = int(g["cfg"][1])
    lut = orc.c_lut_build_vqpy(g["x"][-1], g["codebook"], M)
    assert np.array_equal(lut.view(np.uint16), g["lut_last"].view(np.uint16))
    assert np.array_equal(orc.np_compute_lut(g["x"][-1], g["codebook"]).view(np.uint16), g["lut_last"].view(np.uint16))


@pytest.mark.parametrize("name", VQ_CASES)
def test_vqpy_lut_gemm_bit_exact(orc, golden_dir, name):
    g = _load(golden_dir, name)
    M, B = int(g["cfg"][1]), int(g["cfg"][5])
    out = np.stack([orc.c_lut_gemv_f64(orc.c_lut_build_vqpy(g["x"][b], g["codebook"], M), g["codeword"]) for b in range(B)])
    assert np.array_equal(out, g["out_lut_gemm"])
    assert np.array_equal(orc.np_compute_lut_gemm(g["x"], g["codebook"], g["codeword"]), g["out_lut_gemm"])


@pytest.mark.parametrize("name", VQ_CASES)
def test_vqpy_dequantize_bit_exact(orc, golden_dir, name):
    g = _load(golden_dir, name)
    M = int(g["cfg"][1])
    W = orc.c_dequantize(g["codebook"], g["codeword"], None, M)
    assert np.array_equal(W, g["weight_mat"].view(np.uint16))
    assert np.array_equal(orc.np_dequantize(g["codebook"], g["codeword"]).view(np.uint16), g["weight_mat"].view(np.uint16))
    # compute_fp_gemm (vq.py:304-307) from the dequantized weights
    fp = g["x"].astype(np.float32) @ W.view(np.float16).astype(np.float32).T
    np.testing.assert_allclose(fp, g["out_fp_gemm"], rtol=1e-5, atol=1e-4)


@pytest.mark.parametrize("name", AQLM_CASES)
def test_aqlm_pack_unpack(orc, golden_dir, name):
    g = _load(golden_dir, name)
    assert np.array_equal(orc.np_unpack_int_data(g["codes"]), g["idx"])
    assert np.array_equal(orc.np_pack_int_data(g["idx"]), g["codes"])
    codes = np.ascontiguousarray(g["codes"])           # NpzFile returns a fresh array per access: keep it alive
    out = np.empty(codes.size, dtype=np.uint8)
    orc.lib().orc_unpack_codes(codes.ctypes.data, codes.size, out.ctypes.data)
    assert np.array_equal(out.reshape(codes.shape), g["idx"])


@pytest.mark.parametrize("name", AQLM_CASES)
def test_aqlm_numba_gemv(orc, golden_dir, name):
    g = _load(golden_dir, name)
    I, O, N, B, _, wb = [int(v) for v in g["cfg"]]
    bias = g["bias"] if wb else None
    codes_alt = np.ascontiguousarray(np.transpose(g["codes"], (1, 0, 2))).view(np.uint8)
    ys = np.stack([orc.c_aqlm_gemv_f32(g["x"][b], codes_alt, g["codebooks"], g["scales"], bias) for b in range(B)])
    np.testing.assert_allclose(ys, g["y_numba"], rtol=2e-6, atol=2e-6 * np.abs(g["y_numba"]).max())
    yn = np.stack([orc.np_numba_gemv_lut(g["x"][b], g["codebooks"], codes_alt, g["scales"]) for b in range(B)])
    if wb:
        yn = yn + bias
    np.testing.assert_allclose(yn, g["y_numba"], rtol=1e-4, atol=1e-4 * np.abs(g["y_numba"]).max())


@pytest.mark.parametrize("name", AQLM_CASES)
def test_aqlm_dequantize_weight(orc, golden_dir, name):
    g = _load(golden_dir, name)
    I, O, N = [int(v) for v in g["cfg"][:3]]
    cb16 = g["codebooks"].astype(np.float16).reshape(N, 256, 8)
    W = orc.c_dequantize(cb16, g["idx"], None, I // 8)
    Ws = orc.c_dequantize(cb16, g["idx"], g["scales"].astype(np.float16), I // 8)
    ref, refs = g["weight_f16_noscale"].astype(np.float16), g["weight_f16_scaled"].astype(np.float16)
    if N <= 2:      # one IEEE add: bit-exact class
        assert np.array_equal(W, ref.view(np.uint16))
        assert np.array_equal(Ws, refs.view(np.uint16))
    else:           # embedding_bag sums N>2 terms in its own order: tolerance
        np.testing.assert_allclose(W.view(np.float16).astype(np.float32), ref.astype(np.float32), rtol=2e-3, atol=2e-3)
    w32 = orc.np_aqlm_dequantize_weight(g["idx"], g["codebooks"], g["scales"])
    np.testing.assert_allclose(w32, g["weight_f32"], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("name", AQLM_CASES)
def test_aqlm_exact_and_rowmajor(orc, golden_dir, name):
    """y64 (float64 from raw operands) agrees with the reference's dense check; the timed row-major port agrees with numba."""
    g = _load(golden_dir, name)
    I, O, N, B, _, wb = [int(v) for v in g["cfg"]]
    bias = g["bias"] if wb else None
    x16, cb16, sc16 = g["x"][0].astype(np.float16), g["codebooks"].astype(np.float16), g["scales"].astype(np.float16)
    y64 = orc.c_aqlm_gemv_f64(x16, g["codes"], cb16)
    dense = (g["y_dense"][0] - (bias if wb else 0)) / g["scales"].reshape(-1)
    assert np.abs(y64 - dense).max() / np.abs(dense).max() < 1e-5
    yr = orc.c_aqlm_gemv_rowmajor(x16, g["codes"], cb16, sc16, None if not wb else bias.astype(np.float16))
    yr = yr.view(np.float16).astype(np.float32)
    assert np.abs(yr - g["y_numba"][0]).max() / np.abs(g["y_numba"][0]).max() < 1e-3


def test_simd_lut_scalar_wraparound(orc):
    """bench_avx_lut.cpp:98-107 semantics: int8 LUT values accumulate into int16 with wraparound."""
    rng = np.random.default_rng(0)
    idx = rng.integers(0, 16, size=(100, 16), dtype=np.uint8)
    lut = (np.arange(16) * 10 + 100).astype(np.int64)        # 100..250 stored as int8 (wraps negative), :185
    lut8 = lut.astype(np.uint8).view(np.int8)
    got = orc.c_simd_lut_scalar(idx, lut8, n_runs=37)
    want = (37 * lut8.astype(np.int64)[idx].sum(axis=0)).astype(np.int64)
    want = ((want + 2**15) % 2**16 - 2**15).astype(np.int16)
    assert np.array_equal(got, want)
