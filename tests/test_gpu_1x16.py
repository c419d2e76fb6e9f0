"""GPU parity of the AQLM 1x16 entry points (code1x16.cu) through the C-ABI: dequantized weights bit-exact
against the fixtures the reference's own unpack_int_data + _dequantize_weight produced, GEMV outputs within the north-star
tolerance (max|y - y64| / max|y64| <= 1e-3) of the float64 evaluation of the same fp16 operands."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-3
CASES = ["aqlm_1x16_g8_i256_o48", "aqlm_1x16_g16_i512_o40"]


@pytest.fixture(scope="module")
def ops():
    import rvv_lut_vq_b200 as r
    assert os.path.exists(r._lib.LIB_PATH), "liblutvq.so must be prebuilt in-tree (no fallback)"
    assert torch.cuda.is_available()
    from rvv_lut_vq_b200 import ops
    return ops


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.int16).numpy().view(np.uint16)


def rel_err(got, want):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    return np.abs(got - want).max() / max(np.abs(want).max(), 1e-30)


@pytest.mark.parametrize("name", CASES)
def test_1x16_fixture(ops, orc, golden_dir, name):
    g = dict(np.load(os.path.join(golden_dir, name + ".npz")))
    I, O, gs, B, seed, wb = [int(v) for v in g["cfg"]]
    cb = orc.codebook_1x16(seed, gs)
    codes, cbd, sc = cu(g["codes"]), cu(cb), cu(g["scales"])
    # dequantize: bit-exact, with and without the scale
    assert np.array_equal(bits(ops.code1x16_dequant(codes, cbd, None)), g["weight_f16_noscale"].view(np.uint16))
    assert np.array_equal(bits(ops.code1x16_dequant(codes, cbd, sc)), g["weight_f16_scaled"].view(np.uint16))
    # matmat (batch rows share the codebook fetches) against float64 + the reference's rounding order
    bias = cu(g["bias"]) if wb else None
    y = ops.code1x16_matmat(cu(g["x"]), codes, cbd, sc, bias)
    assert y.shape == (B, O)
    W = g["weight_f16_noscale"]
    for b in range(B):
        acc = orc.c_dense_gemv_f64(g["x"][b], W)
        want = orc.c_scale_bias_epilogue(orc.to_bits(acc.astype(np.float32), 0).reshape(1, -1), g["scales"].reshape(-1), g["bias"] if wb else None, 0).reshape(-1)
        got = y[b].float().cpu().numpy()
        assert rel_err(got, want.view(np.float16).astype(np.float64)) <= TOL
        y64 = acc * g["scales"].reshape(-1).astype(np.float64) + (g["bias"].astype(np.float64) if wb else 0.0)
        assert rel_err(got, y64) <= 2e-3                     # + the fp16 roundings of the epilogue
    # raw matvec launcher (no scale) and the dequant + library GEMM routes
    C = torch.empty(O, dtype=torch.float16, device="cuda")
    ops.code1x16_matvec(codes.squeeze(2), cu(g["x"][0]), C, cbd, False)
    assert rel_err(C.float().cpu().numpy(), orc.c_dense_gemv_f64(g["x"][0], W)) <= TOL
    y2 = ops.code1x16_matmat_dequant(cu(g["x"]), codes, cbd, sc, bias)
    assert rel_err(y2.float().cpu().numpy(), g["y_dense"]) <= 4e-3
    gin = torch.randn((2, O), dtype=torch.float16, device="cuda")
    yt = ops.code1x16_matmat_dequant_transposed(gin, codes, cbd, sc, None)
    want_t = (gin.double().cpu().numpy() * g["scales"].reshape(1, -1).astype(np.float64)) @ W.astype(np.float64)
    assert rel_err(yt.double().cpu().numpy(), want_t) <= 4e-3


@pytest.mark.parametrize("I,O,gs,batch,dtype", [(4096, 4096, 8, 1, "f16"), (4096, 11008, 8, 5, "f16"), (11008, 4096, 8, 2, "f16"),
                                               (4096, 1024, 16, 3, "f16"), (1024, 515, 8, 9, "bf16"), (2048, 33, 16, 1, "bf16")])
def test_1x16_vs_oracle_llm_sizes(ops, orc, I, O, gs, batch, dtype):
    """Llama-2-7B shapes incl. ragged row counts and batches that are not multiples of the 4-row tile; fp16 and bf16."""
    rng = np.random.default_rng(I + O + gs)
    dt = 0 if dtype == "f16" else 1
    tdt = torch.float16 if dt == 0 else torch.bfloat16
    idx = rng.integers(0, 65536, size=(O, I // gs, 1))
    codes = orc.np_pack_int_data(idx, 16)
    cb = torch.from_numpy(rng.standard_normal((1, 65536, 1, gs)).astype(np.float32)).to(tdt)
    x = torch.from_numpy(rng.standard_normal((batch, I)).astype(np.float32)).to(tdt)
    sc = torch.from_numpy(rng.standard_normal((O, 1, 1, 1)).astype(np.float32)).to(tdt)
    W = ops.code1x16_dequant(cu(codes), cb.cuda(), None)
    Wref = cb.reshape(65536, gs)[torch.from_numpy(idx[..., 0])].reshape(O, I)
    assert torch.equal(W.cpu().view(torch.int16), Wref.contiguous().view(torch.int16))
    Ws = ops.code1x16_dequant(cu(codes), cb.cuda(), sc.cuda())
    assert torch.equal(Ws.cpu().view(torch.int16), (Wref * sc.reshape(-1, 1)).contiguous().view(torch.int16))       # one RN multiply
    y = ops.code1x16_matmat(x.cuda(), cu(codes), cb.cuda(), sc.cuda(), None)
    y64 = (x.double() @ Wref.double().T)
    got_unscaled = y.double().cpu() / sc.reshape(1, -1).double()
    assert rel_err(got_unscaled.numpy(), y64.numpy()) <= (TOL if dt == 0 else 8e-3)      # bf16 outputs carry 8 mantissa bits


def test_1x16_errors(ops):
    z16 = torch.zeros((8, 8, 1), dtype=torch.int16, device="cuda")
    with pytest.raises(NotImplementedError):     # group size 4: cuda_kernel.cpp:136-145
        ops.code1x16_matmat(torch.zeros((1, 32), dtype=torch.float16, device="cuda"), z16, torch.zeros((1, 65536, 1, 4), dtype=torch.float16, device="cuda"),
                            torch.ones((8, 1, 1, 1), dtype=torch.float16, device="cuda"))
    with pytest.raises(NotImplementedError):     # fp32 input: cuda_kernel.cpp:9-25
        ops.code1x16_matmat(torch.zeros((1, 64), dtype=torch.float32, device="cuda"), z16, torch.zeros((1, 65536, 1, 8), dtype=torch.float16, device="cuda"),
                            torch.ones((8, 1, 1, 1), dtype=torch.float16, device="cuda"))
    from rvv_lut_vq_b200._lib import LutvqError
    with pytest.raises(LutvqError):              # prob_k not a multiple of 8 codes
        ops.code1x16_matmat(torch.zeros((1, 40), dtype=torch.float16, device="cuda"), torch.zeros((8, 5, 1), dtype=torch.int16, device="cuda"),
                            torch.zeros((1, 65536, 1, 8), dtype=torch.float16, device="cuda"), torch.ones((8, 1, 1, 1), dtype=torch.float16, device="cuda"))


def test_1x16_launch_plan_overrides_are_bit_identical(ops, orc):
    """the kernel's other launch shapes (rows shared by 2 / 4 warps with a shared-memory reduction, more / fewer CTAs per SM; reachable
    through lutvq_debug_set_1x16_plan, which tools/sweep_1x16.py uses) give the same bits as the default plan up to the summation
    order of the row splits: exact equality for split 1, 1e-3 of the layer's range otherwise"""
    from rvv_lut_vq_b200._lib import lib
    rng = np.random.default_rng(3)
    I, O = 4096, 1000
    idx = rng.integers(0, 65536, size=(O, I // 8, 1))
    codes = cu(orc.np_pack_int_data(idx, 16))
    cb = torch.from_numpy(rng.standard_normal((1, 65536, 1, 8)).astype(np.float32)).half().cuda()
    sc = torch.from_numpy(rng.standard_normal((O, 1, 1, 1)).astype(np.float32)).half().cuda()
    x = torch.from_numpy(rng.standard_normal((3, I)).astype(np.float32)).half().cuda()
    want = ops.code1x16_matmat(x, codes, cb, sc, None)
    try:
        for split, per_sm in ((1, 1), (1, 6), (2, 3), (4, 2)):
            lib().lutvq_debug_set_1x16_plan(split, per_sm)
            got = ops.code1x16_matmat(x, codes, cb, sc, None)
            if split == 1:
                assert torch.equal(got.view(torch.int16), want.view(torch.int16))
            else:
                assert rel_err(got.float().cpu().numpy(), want.float().cpu().numpy()) <= TOL
    finally:
        lib().lutvq_debug_set_1x16_plan(0, 0)
