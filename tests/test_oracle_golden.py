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


def test_simd_lut_scalar_matches_reference_build(orc, golden_dir):
    """a13 pinned by the reference itself: fixture produced by bench_avx_lut.cpp -D_SCALAR_ (oracle/_ref/libsimdlut_ref.so,
    oracle/gen_golden_tmac.py): non-zero start values, 300 runs over 500 rows of CB_SIZE 16."""
    g = _load(golden_dir, "simd_lut_ref")
    got = orc.c_simd_lut_scalar(g["idx"], g["lut"], n_runs=int(g["n_runs"][0]), out0=g["out_in"])
    assert np.array_equal(got, g["out"])


# ------------------------------------------------------------------ T-MAC integer-LUT kernels (a14)
def test_tmac_lut_ctor_known_answer_and_fixtures(orc, golden_dir):
    """lut_ctor.cc:38-266 via tmac_oracle.c vs fixtures produced by the reference's compiled code, incl. the known answer
    of its own test_lut_ctor.cc (b[i] = i -> lut_scales 0.929134, lut_biases -496)."""
    g = _load(golden_dir, "tmac_lut_ctor")
    for name in ("kat", "rand32", "rand256", "sat"):
        b = g[f"{name}__b"]
        n = b.shape[0] // 32
        scales = np.array([orc.c_partial_max_g4_int8_k8(0.0, b[32 * k:32 * k + 32]) for k in range(n)], dtype=np.float32)
        assert np.array_equal(scales.view(np.uint32), g[f"{name}__scales"].view(np.uint32))
        for k in range(n):
            qlut, ls, lb = orc.c_lut_ctor_g4_int8(32, b[32 * k:32 * k + 32], scales[k])
            assert np.array_equal(qlut, g[f"{name}__qlut"][8 * k:8 * k + 8])
            assert np.float32(ls).view(np.uint32) == g[f"{name}__lut_scales"][k].view(np.uint32)
            assert np.float32(lb).view(np.uint32) == g[f"{name}__lut_biases"][k].view(np.uint32)
        qs, ls, lb = orc.c_lut_ctor_g4_int8(b.shape[0], b, float(g[f"{name}__scale_single"][0]))
        assert np.array_equal(qs, g[f"{name}__qlut_single"]) and np.float32(lb).view(np.uint32) == g[f"{name}__bias_single"][0].view(np.uint32)
        # the fused preprocessor protocol (kernels.cc:1002-1037)
        q2, s2, b2 = orc.c_tmac_preprocessor_int8(b)
        assert np.array_equal(q2, g[f"{name}__qlut"]) and np.array_equal(s2, g[f"{name}__lut_scales"]) and np.array_equal(b2, g[f"{name}__lut_biases"])
    kat_q = g["kat__qlut"]
    assert kat_q[0].tolist() == [-6, -6, -4, -4, -2, -2, 0, 0, 0, 0, 2, 2, 4, 4, 6, 6]
    assert kat_q[7].tolist() == [-127, -67, -65, -4, -62, -2, 0, 60, -60, 0, 2, 62, 4, 65, 67, 127]
    assert abs(float(g["kat__scales"][0]) - 0.929134) < 1e-6 and float(g["kat__lut_biases"][0]) == -496.0


def test_tmac_tbl_updates_bit_exact(orc, golden_dir):
    """tbl.cc:536-630 (int32) and :323-532 (float: plain, ZeroPoint, FastAggregation) vs the reference's compiled output"""
    g = _load(golden_dir, "tmac_tbl")
    for K, bits in ((16, 2), (8, 4)):
        t = f"i32_k{K}_b{bits}"
        got = orc.c_tbl_g4_int8_int32_update(256, g[f"{t}__c_in"], g[f"{t}__lut"], g[f"{t}__a"], K)
        assert np.array_equal(got, g[f"{t}__c_out"])
    for t in ("f_k16_b2", "f_k8_b4", "f_k16_b4_zp", "f_k16_b2_fa"):
        K, bits, ak, fa, zp = [int(v) for v in g[f"{t}__cfg"]]
        args = (g[f"{t}__lut"], g[f"{t}__a"], g[f"{t}__scales"], g[f"{t}__lut_scales"], g[f"{t}__lut_biases"], K, bits, ak, fa, zp)
        c1 = orc.c_tbl_g4_int8_float_update(256, g[f"{t}__c_in"], *args)
        assert np.array_equal(c1.view(np.uint32), g[f"{t}__c_out"].view(np.uint32)), t
        c2 = orc.c_tbl_g4_int8_float_update(256, c1, *args)
        assert np.array_equal(c2.view(np.uint32), g[f"{t}__c_out2"].view(np.uint32)), t


def test_tmac_qgemm_bit_exact(orc, golden_dir):
    """whole bit-serial GEMV in the generated kernels' call structure (T-MAC deploy/tuned/kernels.cc:827-960)"""
    g = _load(golden_dir, "tmac_qgemm")
    M, K, bits, bm = [int(v) for v in g["cfg"]]
    q, ls, lb = orc.c_tmac_preprocessor_int8(g["B"])
    assert np.array_equal(q, g["qlut"]) and np.array_equal(ls, g["lut_scales"]) and np.array_equal(lb, g["lut_biases"])
    C = orc.c_tmac_qgemm_lut_int8(M, K, bits, bm, g["A"], q, g["scales"], ls, lb)
    assert np.array_equal(C.view(np.uint32), g["C"].view(np.uint32))


def test_tmac_oracle_vs_reference_build_random(orc):
    """when oracle/_ref exists (build container): fresh random cases straight against the reference's compiled code"""
    import ctypes
    ref = orc.ref_tmac()
    if ref is None:
        pytest.skip("oracle/_ref/libtmac_ref.so not built (no /root/reference here)")
    rng = np.random.default_rng(7)
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    for trial in range(20):
        lut = rng.integers(-128, 128, size=(8, 16), dtype=np.int8)
        a = rng.integers(0, 256, size=(128 * 8,), dtype=np.uint8)
        sc = rng.standard_normal(64).astype(np.float32)
        ls, lb = rng.standard_normal(1).astype(np.float32), rng.standard_normal(1).astype(np.float32)
        c = rng.standard_normal(256).astype(np.float32)
        want = c.copy()
        ref.tbl_g4_int8_float_update_strue_k8_b4_ak8_fafalse_zfalse_osfalse(256, P(want), P(lut), P(a), P(sc), P(ls), P(lb))
        got = orc.c_tbl_g4_int8_float_update(256, c, lut, a, sc, ls, lb, 8, 4, 8)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        b = (rng.standard_normal(32) * 10 ** rng.uniform(-3, 3)).astype(np.float32)
        s = np.zeros(1, dtype=np.float32)
        ref.partial_max_g4_int8_k8(P(s), P(b))
        assert np.float32(orc.c_partial_max_g4_int8_k8(0.0, b)).view(np.uint32) == s.view(np.uint32)[0]
        q = np.zeros((8, 16), dtype=np.int8)
        s2, b2 = s.copy(), np.zeros(1, dtype=np.float32)
        ref.lut_ctor_g4_int8_k0_b4(32, P(q), P(b), P(s2), P(b2))
        gq, gs, gb = orc.c_lut_ctor_g4_int8(32, b, float(s[0]))
        assert np.array_equal(gq, q) and np.float32(gb).view(np.uint32) == b2.view(np.uint32)[0]


@pytest.mark.parametrize("name", ["aqlm_1x16_g8_i256_o48", "aqlm_1x16_g16_i512_o40"])
def test_aqlm_1x16_dequantize_bit_exact(orc, golden_dir, name):
    """1x16 (a12): the oracle's gather (* scale) vs the reference's unpack_int_data + _dequantize_weight outputs, bit for bit;
    the float64 dense product of those weights vs the reference's F.linear check."""
    g = _load(golden_dir, name)
    I, O, gs, B, seed, wb = [int(v) for v in g["cfg"]]
    cb = orc.codebook_1x16(seed, gs)
    assert int(cb.view(np.uint16).astype(np.uint64).sum()) == int(g["codebook_checksum"][0]), "codebook regeneration drifted"
    assert np.array_equal(orc.np_unpack_int_data(g["codes"], 16)[..., 0], g["idx"][..., 0].astype(np.int64))
    assert np.array_equal(orc.np_pack_int_data(g["idx"].astype(np.int64), 16), g["codes"])
    W = orc.np_aqlm_1x16_dequantize(g["codes"], cb, None)
    Ws = orc.np_aqlm_1x16_dequantize(g["codes"], cb, g["scales"])
    assert np.array_equal(W.view(np.uint16), g["weight_f16_noscale"].view(np.uint16))
    assert np.array_equal(Ws.view(np.uint16), g["weight_f16_scaled"].view(np.uint16))
    y64 = np.stack([orc.c_dense_gemv_f64(g["x"][b], W) for b in range(B)]) * g["scales"].reshape(1, -1).astype(np.float64)
    if wb:
        y64 = y64 + g["bias"].astype(np.float64)
    assert np.abs(y64 - g["y_dense"]).max() / np.abs(g["y_dense"]).max() < 1e-5


# ---- PQ4 fast-scan: pq4_oracle.c vs fixtures produced by the reference's own compiled code ----
PQ4_CASES = ["q1_b32_m32", "q1_b64_m16", "q1_b96_m8", "q1_b128_m64", "q1_b160_m600", "q2_b32_m7", "q2_b64_m16_s3", "q3_b32_m64_s200", "q4_b32_m20"]


@pytest.mark.parametrize("name", PQ4_CASES)
def test_pq4_oracle_vs_reference_fixture(orc, golden_dir, name):
    """pq4_fast_scan.cpp:45-137,253-264 and pq4_fast_scan_search_1.cpp:30-222 (gen_golden_pq4.py ran oracle/_ref/libpq4_ref.so)."""
    g = _load(golden_dir, "pq4_" + name)
    nq, bbs, M, ntotal, scale = [int(v) for v in g["cfg"]]
    nsq, nb = orc.pq4_nsq(M), -(-ntotal // bbs) * bbs
    blocks = orc.c_pq4_pack_codes(g["codes"], M, nb, bbs)
    assert np.array_equal(blocks, g["blocks"])
    cut = ntotal // 3 + 1
    br = orc.aligned_u8(nb * nsq // 2)
    orc.c_pq4_pack_codes_range(g["codes"][:cut], M, 0, cut, bbs, br)
    orc.c_pq4_pack_codes_range(g["codes"][cut:], M, cut, ntotal, bbs, br)
    assert np.array_equal(br, g["blocks"])
    lp = orc.c_pq4_pack_LUT(g["lut"])
    assert np.array_equal(lp, g["lut_packed"])
    accu = orc.c_pq4_accumulate_loop(nq, nb, bbs, nsq, blocks, lp, scale)
    assert np.array_equal(accu, g["accu"])
    assert np.array_equal(accu[:, :ntotal], orc.np_pq4_distances(g["codes"], g["lut"], M, scale))      # the closed form on unpacked codes
    got = [orc.c_pq4_get_packed_element(blocks, bbs, nsq, int(v), int(s)) for v, s in zip(g["elem_v"], g["elem_sq"])]
    assert got == g["elems"].tolist()


def test_pq4_oracle_packers_and_errors(orc, golden_dir):
    g = _load(golden_dir, "pq4_packers")
    nsq, bbs, qbs = int(g["nsq"][0]), int(g["bbs"][0]), int(g["qbs"][0])
    assert np.array_equal(orc.c_pq4_pack_LUT_qbs(qbs, g["lut"]), g["lut_qbs"])
    assert np.array_equal(orc.c_pq4_pack_LUT_qbs(qbs, g["lut"], g["q_map"]), g["lut_qbs_qmap"])
    block = orc.aligned_u8(g["block"].size)                      # CodePackerPQ4::pack_1 = set_packed_element per nibble
    for flat, off in zip(g["flat"], g["offsets"]):
        for i, c in enumerate(flat):
            orc.c_pq4_set_packed_element(block, int(c) & 15, bbs, nsq, int(off), 2 * i)
            orc.c_pq4_set_packed_element(block, int(c) >> 4, bbs, nsq, int(off), 2 * i + 1)
    assert np.array_equal(block, g["block"])
    with pytest.raises(ValueError):                              # nq = 5 is not instantiated (pq4_fast_scan_search_1.cpp:156-168)
        orc.c_pq4_accumulate_loop(5, 32, 32, 2, orc.aligned_u8(32), orc.aligned_u8(5 * 32))
    with pytest.raises(ValueError):                              # nb % bbs
        orc.c_pq4_accumulate_loop(1, 32, 64, 2, orc.aligned_u8(64), orc.aligned_u8(32))
    with pytest.raises(ValueError):                              # bbs % 32 (pq4_fast_scan.cpp:53)
        orc.c_pq4_pack_codes(np.zeros((4, 1), np.uint8), 2, 48, 48)
