#!/usr/bin/env python
"""Golden vectors for the T-MAC integer-LUT rows -- TEST INFRASTRUCTURE.

Runs the REFERENCE'S OWN compiled code (oracle/_ref/libtmac_ref.so = benchmark/kernel/bs_lut_gemm/{tbl,lut_ctor}.cc and
oracle/_ref/libsimdlut_ref.so = benchmark/kernel/simd_lut/bench_avx_lut.cpp -D_SCALAR_, both built from /root/reference in
place by `make -C oracle ref`) on seeded inputs and freezes inputs + outputs under tests/golden/tmac_*.npz /
simd_lut_ref.npz.  Includes the deterministic known answer of the reference's test_lut_ctor.cc (b[i] = i).
Only runnable in the build container; the fixtures travel."""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
P = ctypes.c_void_p


def _p(a):
    return a.ctypes.data_as(P)


def gen_tmac(out_dir: str) -> None:
    import oracle as orc
    orc.build()
    ref = orc.ref_tmac()
    assert ref is not None, "oracle/_ref/libtmac_ref.so missing: make -C oracle ref"
    rng = np.random.default_rng(2024)

    # ---- lut_ctor + partial_max: the known answer of test_lut_ctor.cc, and random activations ----
    cases = {}
    for name, b in (("kat", np.arange(32, dtype=np.float32)),
                    ("rand32", rng.standard_normal(32).astype(np.float32)),
                    ("rand256", (rng.standard_normal(256) * 3).astype(np.float32)),
                    ("sat", np.concatenate([np.full(4, 1000.0), rng.standard_normal(28)]).astype(np.float32))):
        act_k = b.shape[0]
        # preprocessor protocol (kernels.cc:1002-1037): one scale per 32 activations
        scales = np.zeros(act_k // 32, dtype=np.float32)
        for kk in range(act_k // 32):
            ref.partial_max_reset(_p(scales[kk:]))
            ref.partial_max_g4_int8_k8(_p(scales[kk:]), _p(b[kk * 32:]))
        qlut = np.zeros((act_k // 4, 16), dtype=np.int8)
        ls, lb = scales.copy(), np.zeros(act_k // 32, dtype=np.float32)
        for kk in range(act_k // 32):
            ref.lut_ctor_g4_int8_k0_b4(32, _p(qlut[kk * 8:]), _p(b[kk * 32:]), _p(ls[kk:]), _p(lb[kk:]))
        cases[name] = dict(b=b, scales=scales, qlut=qlut, lut_scales=ls, lut_biases=lb)
        # one multi-block call (act_k > 32 with a single scale), as the impl's loop allows (lut_ctor.cc:127)
        q2 = np.zeros((act_k // 4, 16), dtype=np.int8)
        s2, b2 = np.array([scales.max()], dtype=np.float32), np.zeros(1, dtype=np.float32)
        ref.lut_ctor_g4_int8_k0_b4(act_k, _p(q2), _p(b), _p(s2), _p(b2))
        cases[name].update(qlut_single=q2, scale_single=s2, bias_single=b2)
    kat = cases["kat"]
    assert abs(float(kat["scales"][0]) - 0.929134) < 1e-6 and float(kat["lut_biases"][0]) == -496.0
    assert kat["qlut"][0].tolist() == [-6, -6, -4, -4, -2, -2, 0, 0, 0, 0, 2, 2, 4, 4, 6, 6]
    assert kat["qlut"][7].tolist() == [-127, -67, -65, -4, -62, -2, 0, 60, -60, 0, 2, 62, 4, 65, 67, 127]
    flat = {}
    for n, d in cases.items():
        for k, v in d.items():
            flat[f"{n}__{k}"] = v
    np.savez_compressed(os.path.join(out_dir, "tmac_lut_ctor.npz"), **flat)

    # ---- tbl updates ----
    out = {}
    m = 256

    def tbl_inputs(K, bits, n_sc):
        lut = rng.integers(-127, 128, size=(K, 16), dtype=np.int8)
        a = rng.integers(0, 256, size=(m // 2 * K,), dtype=np.uint8)
        scales = rng.standard_normal(n_sc).astype(np.float32)
        return lut, a, scales

    for (K, bits) in ((16, 2), (8, 4)):
        lut, a, _ = tbl_inputs(K, bits, 1)
        c = rng.integers(-1000, 1000, size=m, dtype=np.int32)
        c_out = c.copy()
        getattr(ref, f"tbl_g4_int8_int32_update_strue_k{K}_b{bits}_ak{K}_fafalse_zfalse_osfalse")(m, _p(c_out), _p(lut), _p(a))
        out.update({f"i32_k{K}_b{bits}__lut": lut, f"i32_k{K}_b{bits}__a": a, f"i32_k{K}_b{bits}__c_in": c, f"i32_k{K}_b{bits}__c_out": c_out})
    for tag, fn, K, bits, ak, fa, zp in (
            ("f_k16_b2", "tbl_g4_int8_float_update_strue_k16_b2_ak16_fafalse_zfalse_osfalse", 16, 2, 16, 0, 0),
            ("f_k8_b4", "tbl_g4_int8_float_update_strue_k8_b4_ak8_fafalse_zfalse_osfalse", 8, 4, 8, 0, 0),
            ("f_k16_b4_zp", "tbl_g4_int8_float_update_strue_k16_b4_ak16_fafalse_ztrue_osfalse", 16, 4, 16, 0, 1),
            ("f_k16_b2_fa", "tbl_g4_int8_float_update_strue_k16_b2_ak16_fatrue_zfalse_osfalse", 16, 2, 16, 1, 0)):
        n_sc = (m // bits) * (2 if zp else 1)
        lut, a, scales = tbl_inputs(K, bits, n_sc)
        ls = np.abs(rng.standard_normal(K // ak)).astype(np.float32) + 0.1
        lb = rng.standard_normal(K // ak).astype(np.float32)
        c = rng.standard_normal(m).astype(np.float32)
        c_out = c.copy()
        getattr(ref, fn)(m, _p(c_out), _p(lut), _p(a), _p(scales), _p(ls), _p(lb))
        # a second accumulating call on the result (test_tbl.cc calls the update twice)
        c_out2 = c_out.copy()
        getattr(ref, fn)(m, _p(c_out2), _p(lut), _p(a), _p(scales), _p(ls), _p(lb))
        out.update({f"{tag}__cfg": np.array([K, bits, ak, fa, zp]), f"{tag}__lut": lut, f"{tag}__a": a, f"{tag}__scales": scales,
                    f"{tag}__lut_scales": ls, f"{tag}__lut_biases": lb, f"{tag}__c_in": c, f"{tag}__c_out": c_out, f"{tag}__c_out2": c_out2})
    np.savez_compressed(os.path.join(out_dir, "tmac_tbl.npz"), **out)

    # ---- whole GEMV in the generated kernels' call structure (kernels.cc:827-960, 1002-1037), reference functions only ----
    M, K, bits, bm = 128, 256, 4, 256
    B = rng.standard_normal(K).astype(np.float32)
    tiles, kt_n = M * bits // bm, K // 32
    A = rng.integers(0, 256, size=(tiles, kt_n, bm // 2 * 8), dtype=np.uint8)
    Sc = (rng.standard_normal((tiles, kt_n, bm // bits)) * 0.1).astype(np.float32)
    ls, lb = np.zeros(kt_n, dtype=np.float32), np.zeros(kt_n, dtype=np.float32)
    qlut = np.zeros((K // 4, 16), dtype=np.int8)
    for kk in range(kt_n):
        ref.partial_max_reset(_p(ls[kk:]))
        ref.partial_max_g4_int8_k8(_p(ls[kk:]), _p(B[kk * 32:]))
    for kk in range(kt_n):
        ref.lut_ctor_g4_int8_k0_b4(32, _p(qlut[kk * 8:]), _p(B[kk * 32:]), _p(ls[kk:]), _p(lb[kk:]))
    C = np.zeros(M, dtype=np.float32)
    for t in range(tiles):
        cb = np.zeros(bm, dtype=np.float32)
        ref.tbl_float_reset(bm, _p(cb))
        for kt in range(kt_n):
            ref.tbl_g4_int8_float_update_strue_k8_b4_ak8_fafalse_zfalse_osfalse(bm, _p(cb), _p(qlut[kt * 8:]), _p(A[t, kt]), _p(Sc[t, kt]), _p(ls[kt:]), _p(lb[kt:]))
        cb = cb.reshape(bm // bits // 8, bits, 8)
        acc = cb[:, 0] * np.float32(0.5)
        acc = acc + cb[:, 1]
        acc = acc + cb[:, 2] * np.float32(2.0)
        acc = acc + cb[:, 3] * np.float32(4.0)
        C[t * (bm // bits):(t + 1) * (bm // bits)] = acc.reshape(-1)
    np.savez_compressed(os.path.join(out_dir, "tmac_qgemm.npz"), cfg=np.array([M, K, bits, bm]), B=B, A=A, scales=Sc, qlut=qlut,
                        lut_scales=ls, lut_biases=lb, C=C)

    # ---- a13: the reference's scalar lookup_and_add_simd (CB_SIZE 16) ----
    sl = orc.ref_simd_lut()
    assert sl is not None and sl.ref_simd_lut_cb_size() == 16
    idx = rng.integers(0, 16, size=(500, 16), dtype=np.uint8)
    lut = rng.integers(-128, 128, size=16, dtype=np.int8)
    o = rng.integers(-100, 100, size=16).astype(np.int16)
    o_out = o.copy()
    sl.ref_lookup_and_add_scalar(_p(idx), _p(lut), _p(o_out), ctypes.c_uint64(300), ctypes.c_uint64(500))
    np.savez_compressed(os.path.join(out_dir, "simd_lut_ref.npz"), idx=idx, lut=lut, out_in=o, out=o_out, n_runs=np.array([300]))
    print("[tmac] fixtures written: tmac_lut_ctor.npz tmac_tbl.npz tmac_qgemm.npz simd_lut_ref.npz")


if __name__ == "__main__":
    gen_tmac(os.path.join(os.path.dirname(HERE), "tests", "golden"))
