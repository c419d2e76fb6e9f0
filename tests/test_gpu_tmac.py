"""GPU parity of the T-MAC integer-LUT entry points (tmac.cu) through the C-ABI: bit-exact against the
fixtures the reference's own compiled code produced (tests/golden/tmac_*.npz, oracle/gen_golden_tmac.py) and against
the C oracle (oracle/tmac_oracle.c) on seeded random inputs at larger sizes."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    import rvv_lut_vq_b200 as r
    assert os.path.exists(r._lib.LIB_PATH), "liblutvq.so must be prebuilt in-tree (no fallback)"
    assert torch.cuda.is_available()
    from rvv_lut_vq_b200 import ops
    return ops


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def same_bits(got: torch.Tensor, want: np.ndarray) -> bool:
    g = got.detach().cpu().numpy()
    return g.shape == want.shape and np.array_equal(g.view(np.uint8), np.ascontiguousarray(want).view(np.uint8))


def test_tmac_lut_ctor_fixtures_bit_exact(ops, golden_dir):
    """lut_ctor.cc:38-266 incl. the known answer of test_lut_ctor.cc (b[i] = i -> lut_scales 0.929134, lut_biases -496)."""
    g = _load(golden_dir, "tmac_lut_ctor")
    for name in ("kat", "rand32", "rand256", "sat"):
        b = g[f"{name}__b"]
        n = b.shape[0] // 32
        bd = cu(b)
        # preprocessor protocol, primitive by primitive: reset + partial_max + lut_ctor(32) per 32 activations
        scales = torch.full((n,), 7.0, dtype=torch.float32, device="cuda")
        for k in range(n):
            ops.partial_max_reset(scales[k:k + 1])
            ops.partial_max_g4_int8_k8(scales[k:k + 1], bd[32 * k:32 * k + 32])
        assert same_bits(scales, g[f"{name}__scales"])
        for k in range(n):
            q, ls, lb = ops.lut_ctor_g4_int8(32, bd[32 * k:32 * k + 32], scales[k:k + 1].clone())
            assert same_bits(q, g[f"{name}__qlut"][8 * k:8 * k + 8])
            assert same_bits(ls, g[f"{name}__lut_scales"][k:k + 1]) and same_bits(lb, g[f"{name}__lut_biases"][k:k + 1])
        # one multi-block call with a single scale (lut_ctor.cc:127 loop)
        q, ls, lb = ops.lut_ctor_g4_int8(b.shape[0], bd, cu(g[f"{name}__scale_single"]))
        assert same_bits(q, g[f"{name}__qlut_single"]) and same_bits(lb, g[f"{name}__bias_single"])
        # the fused preprocessor
        q2, s2, b2 = ops.tmac_preprocessor_int8(bd)
        assert same_bits(q2, g[f"{name}__qlut"]) and same_bits(s2, g[f"{name}__lut_scales"]) and same_bits(b2, g[f"{name}__lut_biases"])
    kat = ops.tmac_preprocessor_int8(cu(np.arange(32, dtype=np.float32)))
    assert abs(float(kat[1][0]) - 0.929134) < 1e-6 and float(kat[2][0]) == -496.0
    assert kat[0][0].tolist() == [-6, -6, -4, -4, -2, -2, 0, 0, 0, 0, 2, 2, 4, 4, 6, 6]
    assert kat[0][7].tolist() == [-127, -67, -65, -4, -62, -2, 0, 60, -60, 0, 2, 62, 4, 65, 67, 127]


def test_tmac_partial_max_keeps_larger_scale(ops):
    b = cu(np.full(32, 0.5, dtype=np.float32))
    s = cu(np.array([3.0], dtype=np.float32))
    ops.partial_max_g4_int8_k8(s, b)                       # std::max(*lut_scales, 2 / 127) keeps 3
    assert float(s[0]) == 3.0


def test_tmac_tbl_updates_fixtures_bit_exact(ops, golden_dir):
    """tbl.cc:323-630 on the fixtures of the reference's AVX2 build: int32, float, zero-point and fast-aggregation."""
    g = _load(golden_dir, "tmac_tbl")
    for K, bits in ((16, 2), (8, 4)):
        t = f"i32_k{K}_b{bits}"
        c = cu(g[f"{t}__c_in"])
        ops.tbl_g4_int8_int32_update(256, c, cu(g[f"{t}__lut"]), cu(g[f"{t}__a"]), K)
        assert same_bits(c, g[f"{t}__c_out"])
    for t in ("f_k16_b2", "f_k8_b4", "f_k16_b4_zp", "f_k16_b2_fa"):
        K, bits, ak, fa, zp = [int(v) for v in g[f"{t}__cfg"]]
        args = (cu(g[f"{t}__lut"]), cu(g[f"{t}__a"]), cu(g[f"{t}__scales"]), cu(g[f"{t}__lut_scales"]), cu(g[f"{t}__lut_biases"]), K, bits, ak, bool(fa), bool(zp))
        c = cu(g[f"{t}__c_in"])
        ops.tbl_g4_int8_float_update(256, c, *args)
        assert same_bits(c, g[f"{t}__c_out"]), t
        ops.tbl_g4_int8_float_update(256, c, *args)        # the second, accumulating call of test_tbl.cc
        assert same_bits(c, g[f"{t}__c_out2"]), t
    c = ops.tbl_reset(cu(np.arange(64, dtype=np.float32)))
    assert float(c.abs().max()) == 0.0
    c = ops.tbl_reset(cu(np.arange(64, dtype=np.int32)))
    assert int(c.abs().max()) == 0


def test_tmac_qgemm_fixture_bit_exact(ops, golden_dir):
    g = _load(golden_dir, "tmac_qgemm")
    M, K, bits, bm = [int(v) for v in g["cfg"]]
    q, ls, lb = ops.tmac_preprocessor_int8(cu(g["B"]))
    assert same_bits(q, g["qlut"]) and same_bits(ls, g["lut_scales"]) and same_bits(lb, g["lut_biases"])
    C = ops.tmac_qgemm_lut_int8(M, K, bits, bm, cu(g["A"]), q, cu(g["scales"]), ls, lb)
    assert same_bits(C, g["C"])


@pytest.mark.parametrize("M,K,bits,bm", [(4096, 4096, 4, 256), (1024, 2048, 2, 128), (384, 1024, 3, 96), (512, 512, 1, 64), (11008, 4096, 4, 512)])
def test_tmac_qgemm_vs_oracle_llm_sizes(ops, orc, M, K, bits, bm):
    """Whole bit-serial GEMV at LLM sizes against the C oracle: qlut / scales / biases and C bit for bit."""
    rng = np.random.default_rng(M + K + bits)
    B = rng.standard_normal(K).astype(np.float32)
    tiles, kt_n = M * bits // bm, K // 32
    A = rng.integers(0, 256, size=(tiles, kt_n, bm // 2 * 8), dtype=np.uint8)
    Sc = (rng.standard_normal((tiles, kt_n, bm // bits)) * 0.1).astype(np.float32)
    q0, ls0, lb0 = orc.c_tmac_preprocessor_int8(B)
    want = orc.c_tmac_qgemm_lut_int8(M, K, bits, bm, A, q0, Sc, ls0, lb0)
    q, ls, lb = ops.tmac_preprocessor_int8(cu(B))
    assert same_bits(q, np.asarray(q0).reshape(-1, 16)) and same_bits(ls, ls0) and same_bits(lb, lb0)
    C = ops.tmac_qgemm_lut_int8(M, K, bits, bm, cu(A), q, cu(Sc), ls, lb)
    assert same_bits(C, want)


def test_tmac_int32_wraparound_and_large_k(ops, orc):
    """int16 wraparound of the k-sum (SignedWideningAdder, tbl.cc:216-235): K = 512 tables of +-127 overflow int16."""
    rng = np.random.default_rng(5)
    m, K = 512, 512
    lut = np.where(rng.random((K, 16)) < 0.5, -127, 127).astype(np.int8)
    lut[:, 0] = 127
    a = rng.integers(0, 256, size=m // 2 * K, dtype=np.uint8)
    a[:K * 16] = 0                                          # block 0: every lookup hits entry 0 = 127 -> 65024 wraps to -512
    c0 = rng.integers(-5, 5, size=m, dtype=np.int32)
    want = orc.c_tbl_g4_int8_int32_update(m, c0, lut, a, K)
    got = ops.tbl_g4_int8_int32_update(m, cu(c0), cu(lut), cu(a), K)
    assert same_bits(got, want)
    assert int(want[0] - c0[0]) == ((127 * K + 32768) % 65536) - 32768


def test_tmac_errors(ops):
    from rvv_lut_vq_b200._lib import LutvqError
    z = torch.zeros(64, dtype=torch.int32, device="cuda")
    with pytest.raises(LutvqError):
        ops.tbl_g4_int8_int32_update(48, z, torch.zeros(256, dtype=torch.int8, device="cuda"), torch.zeros(1024, dtype=torch.uint8, device="cuda"), 16)
    with pytest.raises(LutvqError):          # no kernel for this tiling: LUTVQ_EUNSUPPORTED, like the reference's -1
        ops.tmac_qgemm_lut_int8(100, 64, 4, 256, torch.zeros(1, dtype=torch.uint8, device="cuda"), torch.zeros(256, dtype=torch.int8, device="cuda"),
                                torch.zeros(1, device="cuda"), torch.zeros(2, device="cuda"), torch.zeros(2, device="cuda"))
