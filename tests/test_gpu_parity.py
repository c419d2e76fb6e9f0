"""GPU parity: the CUDA path (through the C-ABI / the reference-mirroring ops) against the CPU oracle and
the committed golden vectors.  Bit-exact for index / dequantize / golden-model LUT work; stated tolerances
for fp16/fp32 accumulation (max|y - y64| / max|y64| <= 1e-3)."""
import ctypes
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL_GEMV = 1e-3          # north_star: max rel err 1e-3 vs the exact (float64) evaluation
VQ_CASES = ["vq_d128_m16_n2_k256", "vq_d64_m8_n1_k256", "vq_d64_m8_n4_k256", "vq_d96_m6_n2_k256_d16"]
AQLM_2X8 = ["aqlm_2x8_i256_o96", "aqlm_2x8_i1024_o64"]


@pytest.fixture(scope="module")
def pkg():
    import rvv_lut_vq_b200 as r
    assert os.path.exists(r._lib.LIB_PATH), "liblutvq.so must be prebuilt in-tree (no fallback)"
    assert torch.cuda.is_available()
    return r


@pytest.fixture(scope="module")
def ops(pkg):
    from rvv_lut_vq_b200 import ops
    return ops


def cu(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


def bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.int16).numpy().view(np.uint16)


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name + ".npz")))


def expected_epilogue(orc, acc64, scales16, bias16, dtype=0):
    """reference rounding order on the exact accumulator: round -> *scale -> +bias (cuda_kernel.cpp:95-111)"""
    y = orc.to_bits(acc64.astype(np.float32), dtype)
    return orc.c_scale_bias_epilogue(y.reshape(1, -1), scales16, bias16, dtype).reshape(-1)


def rel_err(got, want):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    return np.abs(got - want).max() / max(np.abs(want).max(), 1e-30)


# ------------------------------------------------------------------ golden model (vq.py), bit-exact
@pytest.mark.parametrize("name", VQ_CASES)
def test_vqpy_golden_bit_exact(pkg, ops, golden_dir, name):
    g = _load(golden_dir, name)
    D, M, N, K, d_out, B, _ = [int(v) for v in g["cfg"]]
    cb, cw = cu(g["codebook"]), cu(g["codeword"])
    outs = []
    for b in range(B):
        lut = ops.lut_build(cu(g["x"][b]), cb, M)                       # VQ.compute_lut
        outs.append(ops.lut_gemv(cw, lut).cpu().numpy())                # VQ.compute_lut_gemm (float64)
    assert np.array_equal(bits(lut), g["lut_last"].view(np.uint16)), "LUT entries differ from vq.py"
    assert np.array_equal(np.stack(outs), g["out_lut_gemm"]), "LUT-GEMV differs from vq.py"
    W = ops.vq_dequantize(cb, cw, M)                                    # VQ.dequantize
    assert np.array_equal(bits(W), g["weight_mat"].view(np.uint16)), "dequantized weights differ from vq.py"


def test_vqpy_shared_codebook_matches_oracle(pkg, ops, orc):
    """AQLM is vq.py with the codebook broadcast over subvectors: config-1 shape, seeded."""
    rng = np.random.default_rng(42)
    M, N, K, d, d_out = 512, 2, 256, 8, 256
    cb = rng.standard_normal((N, K, d)).astype(np.float16)
    cw = rng.integers(0, K, size=(d_out, M, N), dtype=np.uint8)
    x = rng.standard_normal(M * d).astype(np.float16)
    lut = ops.lut_build(cu(x), cu(cb), M)
    want_lut = orc.c_lut_build_vqpy(x, cb, M)
    assert np.array_equal(bits(lut), want_lut.view(np.uint16))
    out = ops.lut_gemv(cu(cw), lut).cpu().numpy()
    assert np.array_equal(out, orc.c_lut_gemv_f64(want_lut, cw))
    W = ops.vq_dequantize(cu(cb), cu(cw), M)
    assert np.array_equal(bits(W), orc.c_dequantize(cb, cw, None, M))


# ------------------------------------------------------------------ AQLM goldens (numba / _dequantize_weight)
@pytest.mark.parametrize("name", AQLM_2X8)
def test_aqlm_golden_dequant_bit_exact(pkg, ops, golden_dir, name):
    g = _load(golden_dir, name)
    codes, cb, sc = cu(g["codes"]), cu(g["codebooks"], torch.float16), cu(g["scales"], torch.float16)
    W = ops.code2x8_dequant(codes, cb, None)
    assert np.array_equal(bits(W), g["weight_f16_noscale"].astype(np.float16).view(np.uint16))
    Ws = ops.code2x8_dequant(codes, cb, sc)
    assert np.array_equal(bits(Ws), g["weight_f16_scaled"].astype(np.float16).view(np.uint16))


@pytest.mark.parametrize("name", AQLM_2X8)
def test_aqlm_golden_matmat(pkg, ops, golden_dir, name):
    g = _load(golden_dir, name)
    wb = int(g["cfg"][5])
    x = cu(g["x"], torch.float16)
    codes, cb, sc = cu(g["codes"]), cu(g["codebooks"], torch.float16), cu(g["scales"], torch.float16)
    bias = cu(g["bias"], torch.float16) if wb else None
    y = ops.code2x8_matmat(x, codes, cb, sc, bias).float().cpu().numpy()
    assert y.shape == g["y_numba"].shape
    assert rel_err(y, g["y_numba"]) <= TOL_GEMV                       # vs numba_gemm_lut (fp32) run by the reference
    # unfused pair in numba's own arithmetic: fp32 LUT, fp32 sequential gather -> tight agreement
    from rvv_lut_vq_b200 import _lib
    G = g["codes"].shape[1]
    lut = ops.lut_build(x[0], cb.reshape(2, 256, 8), G, lut_mode=_lib.LUT_FP32)
    acc = ops.lut_gemv(codes.view(torch.uint8), lut, acc_mode=_lib.ACC_F32).cpu().numpy()
    want = (g["y_numba"][0] - (g["bias"] if wb else 0)) / g["scales"].reshape(-1)
    assert rel_err(acc, want) <= 5e-6


# ------------------------------------------------------------------ seeded cases vs the oracle
CASES = [
    # (in, out, scales, bias, batch)        what it covers
    (64, 32, "randn", False, 1),            # single partial k-slice (8 groups), one row set
    (256, 96, "randn", True, 2),            # n_k == 1 direct-epilogue path
    (512, 40, "ones", False, 1),            # out not a multiple of 8: clamped rows
    (2048, 1000, "randn", True, 1),         # out not a multiple of 32, split-k reduce
    (4096, 4096, "randn", False, 1),        # BASELINE configs[0] shape
    (11008, 256, "randn", True, 3),         # 43 k-slices
    (1088, 300, "randn", False, 1),         # G = 136: last k-slice is partial (8 of 32 groups)
]


@pytest.mark.parametrize("I,O,scales,bias,batch", CASES)
def test_matmat_vs_oracle(pkg, ops, orc, I, O, scales, bias, batch):
    from rvv_lut_vq_b200 import synth
    L = synth.make_layer_np(I, O, seed=I + O, scales=scales, bias=bias, batch=batch)
    y = ops.code2x8_matmat(cu(L["x"]), cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"]),
                           None if L["bias"] is None else cu(L["bias"]))
    assert y.shape == (batch, O) and y.dtype == torch.float16
    for b in range(batch):
        y64 = orc.c_aqlm_gemv_f64(L["x"][b], L["codes"], L["codebooks"])
        s64 = L["scales"].reshape(-1).astype(np.float64)
        want64 = y64 * s64 + (0 if L["bias"] is None else L["bias"].astype(np.float64))
        got = y[b].float().cpu().numpy()
        assert rel_err(got, want64) <= TOL_GEMV
        # and against the fp32-accumulate port with the reference's epilogue rounding: at most 2 fp16 ulp apart, an ulp
        # being that of the largest operand of the epilogue (round(acc)*s, bias, result) -- the bias add can cancel,
        # and one ulp of the product is then many ulps of the (small) sum
        port = orc.c_aqlm_gemv_rowmajor(L["x"][b], L["codes"], L["codebooks"], L["scales"], L["bias"])
        portf = port.view(np.float16).astype(np.float64)
        mag = np.maximum(np.abs(portf), np.abs(y64 * s64))
        if L["bias"] is not None:
            mag = np.maximum(mag, np.abs(L["bias"].astype(np.float64)))
        ulp_size = 2.0 ** (np.floor(np.log2(np.maximum(mag, 2.0 ** -14))) - 10)
        # the LUT is 16-bit fixed point with ONE scale per GEMV (gemv_lut.cu): step q = max|C_k|_2 * max|x_g|_2 / 32704 per
        # entry, i.e. an absolute error of about q * sqrt(2G / 12) per row whatever the row's own magnitude -- for a row
        # that cancels to ~0 that exceeds its (tiny) ulp.  Allow 1e-4 of the layer's range on top of the 2 ulp
        # (10x inside the north-star tolerance of 1e-3; the golden model's fp16 LUT, vq.py:285, is ~4x coarser).
        floor_abs = 1e-4 * np.abs(y64).max() * np.abs(s64)
        off = np.maximum(np.abs(got.astype(np.float64) - portf) - floor_abs, 0) / ulp_size
        assert (off <= 2).mean() > 0.999, f"fp16 outputs differ by more than 2 ulp + 1e-4 of range in {(off > 2).mean():.4%} of rows"
        # the quantisation error itself, measured: rms over rows <= 5e-5 of the range
        rms = np.sqrt(np.mean(((got - want64) / np.maximum(np.abs(s64), 1e-3)) ** 2)) / np.abs(y64).max()
        assert rms <= 2e-4, f"rms error {rms:.2e} of the layer's range"


def test_matmat_bf16(pkg, ops, orc):
    from rvv_lut_vq_b200 import synth
    I, O = 1024, 512
    L = synth.make_layer_np(I, O, seed=5, scales="randn", bias=True)
    tb = lambda a: cu(a.astype(np.float32)).to(torch.bfloat16)
    x, cb, sc, bi = tb(L["x"]), tb(L["codebooks"]), tb(L["scales"]), tb(L["bias"])
    y = ops.code2x8_matmat(x, cu(L["codes"]), cb, sc, bi)
    assert y.dtype == torch.bfloat16
    y64 = orc.c_aqlm_gemv_f64(bits(x[0]), L["codes"], bits(cb), dtype=1)
    want = y64 * sc.float().cpu().numpy().reshape(-1) + bi.float().cpu().numpy()
    assert rel_err(y[0].float().cpu().numpy(), want) <= 8e-3          # bf16 output rounding: 2^-8
    W = ops.code2x8_dequant(cu(L["codes"]), cb, sc)
    assert np.array_equal(bits(W), orc.c_dequantize(bits(cb).reshape(2, 256, 8), L["idx"], bits(sc), I // 8, dtype=1))


def test_lut_fp16_flag_matches_vqpy_semantics(pkg, ops, orc):
    """LUTVQ_FLAG_LUT_FP16: fused GEMV with LUT entries rounded to fp16 == vq.py's LUT-GEMV within tolerance."""
    from rvv_lut_vq_b200 import synth
    I, O = 4096, 512
    L = synth.make_layer_np(I, O, seed=9, scales="ones")
    y = ops.code2x8_matmat(cu(L["x"]), cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"]), None, lut_fp16=True)
    lut = orc.c_lut_build_vqpy(L["x"][0], L["codebooks"].reshape(2, 256, 8), I // 8)
    want = orc.c_lut_gemv_f64(lut, L["idx"])
    assert rel_err(y[0].float().cpu().numpy(), want) <= TOL_GEMV


# ------------------------------------------------------------------ full BASELINE sizes
@pytest.mark.parametrize("I,O", [(4096, 4096), (4096, 11008), (11008, 4096)])
def test_full_size_sweep(pkg, ops, orc, I, O):
    """BASELINE configs[1]: the three Llama-2-7B shapes, checked directly against the oracle (it is fast enough)
    plus size-independent properties: determinism, row-permutation equivariance, dequant round trip, zero input."""
    from rvv_lut_vq_b200 import synth
    L = synth.make_layer_np(I, O, seed=1, scales="randn", bias=False)
    x, codes, cb, sc = cu(L["x"]), cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"])
    y = ops.code2x8_matmat(x, codes, cb, sc, None)
    y64 = orc.c_aqlm_gemv_f64(L["x"][0], L["codes"], L["codebooks"]) * L["scales"].reshape(-1).astype(np.float64)
    assert rel_err(y[0].float().cpu().numpy(), y64) <= TOL_GEMV
    # deterministic: fixed-order split-k reduction
    y2 = ops.code2x8_matmat(x, codes, cb, sc, None)
    assert torch.equal(y, y2)
    # permuting output rows IN BLOCKS OF 32 (the kernel's row set; a row's fp32 summation order depends only on
    # its position inside the set) permutes y bit-for-bit: the split-k reduction is integer and order-free
    nblk = O // 32
    bperm = torch.randperm(nblk, device="cuda", generator=torch.Generator(device="cuda").manual_seed(0))
    perm = (bperm[:, None] * 32 + torch.arange(32, device="cuda")[None, :]).reshape(-1)
    yp = ops.code2x8_matmat(x, codes[perm].contiguous(), cb, sc[perm].contiguous(), None)
    assert torch.equal(yp[0], y[0][perm])
    # the prepacked code image and the X_READY overlap mode give bit-identical results
    packed = ops.code2x8_prepack(codes)
    assert torch.equal(ops.code2x8_matmat_packed(x, packed, O, cb, sc, None), y)
    for _ in range(3):
        assert torch.equal(ops.code2x8_matmat_packed(x, packed, O, cb, sc, None, x_ready=True), y)
        assert torch.equal(ops.code2x8_matmat(x, codes, cb, sc, None, x_ready=True), y)
    # dequantize (bit-exact vs oracle on a row sample) and the dense check of benchmark_rq_gemm_gpu.py:99-109
    W = ops.code2x8_dequant(codes, cb, sc)
    rows = np.linspace(0, O - 1, 64).astype(int)
    Wref = orc.c_dequantize(L["codebooks"].reshape(2, 256, 8), L["idx"][rows], L["scales"].reshape(-1)[rows], I // 8)
    assert np.array_equal(bits(W[torch.from_numpy(rows).cuda()]), Wref)
    dense = (x.double() @ W.double().T)[0].cpu().numpy()
    assert rel_err(y[0].double().cpu().numpy(), dense) <= 2e-3        # W carries the extra scale rounding
    # zero activations -> exactly zero
    y0 = ops.code2x8_matmat(torch.zeros_like(x), codes, cb, sc, None)
    assert torch.count_nonzero(y0) == 0


def test_dequant_full_matrix_checksum(pkg, ops, orc):
    """whole 4096x11008 dequant against the oracle via a checksum of row checksums (bit-exact class)"""
    from rvv_lut_vq_b200 import synth
    I, O = 4096, 11008
    L = synth.make_layer_np(I, O, seed=3, scales="randn")
    W = ops.code2x8_dequant(cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"]))
    Wref = orc.c_dequantize(L["codebooks"].reshape(2, 256, 8), L["idx"], L["scales"], I // 8)
    got = bits(W).astype(np.uint64)
    weights = (np.arange(I, dtype=np.uint64) * 2654435761 + 12345) % 1000003
    assert np.array_equal((got * weights).sum(axis=1), (Wref.astype(np.uint64) * weights).sum(axis=1))


# ------------------------------------------------------------------ prefill: dequant -> tensor memory -> tcgen05 GEMM
GEMM_CASES = [
    # (in, out, batch, dtype, bias)
    (256, 128, 16, "f16", False),          # one row tile, 4 k blocks, no split
    (1024, 1000, 17, "f16", True),         # ragged rows (1000 = 7 tiles + 104) and ragged batch (17 -> 32 columns)
    (4096, 4096, 64, "f16", True),         # 32 row tiles -> split-k 4 through the workspace
    (4096, 11008, 256, "f16", False),      # BASELINE configs[2] top batch: 138 CTAs x 80 rows, no split, 3 stages
    (11008, 4096, 128, "f16", False),      # long k (172 blocks), split-k
    (2048, 512, 300, "f16", True),         # batch > 256: chunked launches
    (1024, 768, 48, "bf16", True),
    (4096, 11008, 16, "f16", True),        # 138 CTAs x 80 rows (rows-per-CTA plan), 2 k blocks per pipeline stage
    (11008, 4096, 32, "f16", True),        # split-k 4 x 43 k blocks: odd tail stage with 2 k blocks per stage
    (320, 200, 8, "bf16", False),          # 5 k blocks, ragged rows, one chunk
    (64, 256, 3, "f16", False),            # ONE k block: a single, partial code chunk (fewer chunks than ring slots); no bias: with 64 terms a bias
                                           # cancels enough outputs to near zero that the 2-ulp bar below stops meaning anything (the float64 bar held)
    (512, 40, 129, "f16", False),          # 144 columns -> 1 k block per stage; fewer rows than two row warps
    (1536, 300, 65, "bf16", True),         # 80 columns -> 2 k blocks per stage; 24 k blocks = 6 code chunks > the ring of 4
    (8192, 1024, 40, "f16", True),         # 8 row tiles -> split-k 16 (8 k blocks per CTA = 2 chunks each)
]


@pytest.mark.parametrize("I,O,batch,dt,bias", GEMM_CASES)
def test_matmat_dequant_tensor_core(pkg, ops, orc, I, O, batch, dt, bias):
    """code2x8_matmat_dequant: exact products of the bit-exact dequantized weights, fp32 accumulation in TMEM.
    Checked against float64 of the same operands (tolerance 1e-3 of max |y|, north star) and, tighter, against the
    reference's own recipe (dequant -> F.linear in fp32 -> round -> *scale -> +bias, cuda_kernel.cpp:463-477,95-111)."""
    from rvv_lut_vq_b200 import synth
    tdt = torch.float16 if dt == "f16" else torch.bfloat16
    L = synth.make_layer_np(I, O, seed=7 * I + O + batch, scales="randn", bias=bias, batch=batch)
    conv = (lambda a: cu(a)) if dt == "f16" else (lambda a: cu(a.astype(np.float32)).to(torch.bfloat16))
    x, cb, sc = conv(L["x"]), conv(L["codebooks"]), conv(L["scales"])
    bi = conv(L["bias"]) if bias else None
    codes = cu(L["codes"])
    y = ops.code2x8_matmat_dequant_fused(x, codes, cb, sc, bi)
    assert y.shape == (batch, O) and y.dtype == tdt
    # the oracle's own chain: orc_dequantize (bit-exact weights) -> orc_dense_gemm_f64 (float64 contraction of the same operands)
    dtn = 0 if dt == "f16" else 1
    Wbits = orc.c_dequantize(bits(cb).reshape(2, 256, 8), L["idx"], None, I // 8, dtype=dtn)
    W = ops.code2x8_dequant(codes, cb, None)
    assert np.array_equal(bits(W), Wbits)
    acc64 = torch.from_numpy(orc.c_dense_gemm_f64(bits(x), Wbits, dtn)).cuda()
    want64 = acc64 * sc.reshape(-1).double() + (bi.double() if bias else 0)
    tol = 1e-3 if dt == "f16" else 8e-3
    assert rel_err(y.double().cpu().numpy(), want64.cpu().numpy()) <= tol
    ref = (x.float() @ W.float().T).to(tdt) * sc.reshape(-1)
    if bias:
        ref = ref + bi
    ulp = (y.view(torch.int16).int() - ref.view(torch.int16).int()).abs()
    assert (ulp <= 2).float().mean() > 0.995
    # deterministic (fixed-order split-k) and leaves its workspace clean
    assert torch.equal(ops.code2x8_matmat_dequant_fused(x, codes, cb, sc, bi), y)
    for ws in ops._gemm_workspaces.values():
        assert torch.count_nonzero(ws[:256 + 4 * ((O + 127) // 128)]) == 0     # header + tile counters (partials need no cleaning)
    # the dispatching op (fused below ops.TWO_STEP_MIN_ROWS rows, dequantize-once + library GEMM above) and the two-step route
    # itself: same float64 bar; the two-step route folds the scale into W (one RN multiply per weight, cuda_kernel.cpp:445)
    for fn in (ops.code2x8_matmat_dequant, ops.code2x8_matmat_dequant_two_step):
        y2 = fn(x, codes, cb, sc, bi)
        assert y2.shape == (batch, O) and y2.dtype == tdt
        assert rel_err(y2.double().cpu().numpy(), want64.cpu().numpy()) <= tol


def test_quantized_linear_prefill_dispatch(pkg, orc):
    """batch > 6 -> gemm op (inference.py:95-96); must agree with the gemv op row by row within fp16 rounding"""
    from rvv_lut_vq_b200 import QuantizedLinear, synth
    I, O = 1024, 640
    L = synth.make_layer_np(I, O, seed=77, scales="randn", bias=True, batch=40)
    m = QuantizedLinear(I, O, 8, 1, 2, 8, bias=True, device="cuda", dtype=torch.float16)
    m.load_state_dict({"codes": cu(L["codes"]), "codebooks": cu(L["codebooks"]), "scales": cu(L["scales"]), "bias": cu(L["bias"])})
    x = cu(L["x"])
    y_gemm = m(x)
    y_gemv = torch.cat([m(x[i:i + 4]) for i in range(0, 40, 4)])
    assert rel_err(y_gemm.float().cpu().numpy(), y_gemv.float().cpu().numpy()) <= 1e-3


@pytest.mark.parametrize("I,O,batch,dtype", [(4096, 11008, 2, "f16"), (4096, 4096, 4, "f16"), (11008, 4096, 6, "f16"), (4096, 11008, 11, "bf16"),
                                             (1024, 96, 5, "f16"), (320, 515, 7, "f16"), (1024, 33, 8, "bf16"), (4096, 4096, 8, "f16"),
                                             (64, 8, 2, "f16"), (8192, 1024, 3, "f16")])
def test_multi_row_call_groups_rows_bit_identically(pkg, ops, I, O, batch, dtype):
    """up to 8 rows of one lutvq_code2x8_matmat[_packed] call go out as ONE grouped launch when the workspace holds their split-k
    words (batch 11 = 8 + 3): same bits as one row per call, the workspace comes back all-zero, and x produced by the kernel just
    before the call is honoured (a race would show up as stale rows)."""
    from rvv_lut_vq_b200 import synth
    tdt = torch.float16 if dtype == "f16" else torch.bfloat16
    L = synth.make_layer_np(I, O, seed=I + O + batch, scales="randn", bias=True, batch=batch)
    conv = (lambda a: cu(a)) if dtype == "f16" else (lambda a: cu(a.astype(np.float32)).to(torch.bfloat16))
    x, cb, sc, bi = conv(L["x"]), conv(L["codebooks"]), conv(L["scales"]), conv(L["bias"])
    codes = cu(L["codes"])
    packed = ops.code2x8_prepack(codes)
    rows = torch.cat([ops.code2x8_matmat_packed(x[b:b + 1], packed, O, cb, sc, bi) for b in range(batch)])      # one row per call
    assert ops.BATCH_GROUPED
    got_p = ops.code2x8_matmat_packed(x, packed, O, cb, sc, bi)
    got_r = ops.code2x8_matmat(x, codes, cb, sc, bi)
    torch.cuda.synchronize()
    assert torch.equal(got_p.view(torch.int16), rows.view(torch.int16))
    assert torch.equal(got_r.view(torch.int16), rows.view(torch.int16))
    for ws in ops._batch_workspaces.values():
        assert torch.count_nonzero(ws) == 0
    # producer -> multi-row call -> consumer, many times: x changes every iteration and is written by the kernel right before the call
    acc = torch.zeros((batch, O), dtype=torch.float32, device="cuda")
    want = torch.zeros_like(acc)
    for it in range(20):
        xi = (x.float() * (1.0 + 0.25 * it)).to(tdt)                     # producer kernels
        y = ops.code2x8_matmat_packed(xi, packed, O, cb, sc, bi)
        acc += y.float()                                                 # consumer right behind the last row
    torch.cuda.synchronize()
    for it in range(20):
        xi = (x.float() * (1.0 + 0.25 * it)).to(tdt)
        torch.cuda.synchronize()
        want += torch.cat([ops.code2x8_matmat_packed(xi[b:b + 1], packed, O, cb, sc, bi) for b in range(batch)]).float()
        torch.cuda.synchronize()
    assert torch.equal(acc, want)


# ------------------------------------------------------------------ edge cases / API behaviour
def test_empty_batch_and_shapes(pkg, ops):
    from rvv_lut_vq_b200 import synth
    L = synth.make_layer_np(256, 64, seed=0)
    codes, cb, sc = cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"])
    y = ops.code2x8_matmat(torch.zeros((0, 256), dtype=torch.float16, device="cuda"), codes, cb, sc, None)
    assert y.shape == (0, 64)
    y = ops.code2x8_matmat(torch.randn((2, 3, 256), dtype=torch.float16, device="cuda"), codes, cb, sc, None)
    assert y.shape == (2, 3, 64)                                      # input.shape[:-1] + (out,)  (cuda_kernel.cpp:105-110)
    with pytest.raises(NotImplementedError):
        ops.code2x8_matmat(torch.zeros((1, 256), dtype=torch.float32, device="cuda"), codes, cb, sc, None)


def test_c_abi_direct_and_errors(pkg):
    """call the C-ABI with raw pointers (no torch in the signature); workspace-too-small is EINVAL, not a crash"""
    from rvv_lut_vq_b200 import synth
    lib = pkg.lib()
    L = synth.make_layer_np(1024, 128, seed=2)
    codes, cb, x = cu(L["codes"]), cu(L["codebooks"]), cu(L["x"][0])
    y = torch.empty(128, dtype=torch.float16, device="cuda")
    need = lib.lutvq_workspace_bytes(128, 1024)
    ws = torch.zeros(need, dtype=torch.uint8, device="cuda")
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.lutvq_code2x8_matvec(P(codes), P(x), P(y), P(cb), 128, 1024, 0, P(ws), need, st) == 0
    torch.cuda.synchronize()
    assert lib.lutvq_code2x8_matvec(P(codes), P(x), P(y), P(cb), 128, 1024, 0, P(ws), 16, st) == pkg._lib.LUTVQ_EINVAL
    assert b"workspace" in lib.lutvq_last_error()
    sm = ctypes.c_int()
    assert lib.lutvq_device_info(ctypes.byref(sm), None, None, None) == 0 and sm.value > 0


def test_cuda_graph_and_pdl(pkg, ops, orc):
    """capture a chain of dependent GEMVs (y feeds the next x) in one CUDA graph; replay must equal eager,
    with and without programmatic dependent launch"""
    from rvv_lut_vq_b200 import synth, _lib
    I = O = 1024
    Ls = [synth.make_layer_np(I, O, seed=20 + i, scales="randn") for i in range(4)]
    dev = [{k: cu(v) for k, v in L.items() if v is not None and k != "idx"} for L in Ls]
    for d in dev:
        d["scales"] = d["scales"] * 0.05                              # keep the chain in fp16 range
    x0 = cu(Ls[0]["x"])

    def chain(flags=None):
        h = x0
        for d in dev:
            if flags is None:
                h = ops.code2x8_matmat(h, d["codes"], d["codebooks"], d["scales"], None)
            else:
                h = ops._matmat("lutvq_code2x8_matmat_ex", h, d["codes"], d["codebooks"], d["scales"], None, flags)
        return h

    eager = chain()
    nopdl = chain(_lib.FLAG_NO_PDL)
    assert torch.equal(eager, nopdl)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        chain()                                                       # allocate workspace for this stream
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            out = chain()
    for _ in range(3):
        graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager)


def test_x_ready_overlapping_launches(pkg, ops, orc):
    """LUTVQ_FLAG_X_READY: a burst of independent GEMVs (resident inputs, rotating workspaces) overlaps on the GPU;
    every output must still equal the stream-ordered result, and the workspace must come back all-zero."""
    from rvv_lut_vq_b200 import synth
    I, O = 4096, 1536
    Ls = [synth.make_layer_np(I, O, seed=40 + i, scales="randn") for i in range(6)]
    dev = [{k: cu(v) for k, v in L.items() if v is not None and k != "idx"} for L in Ls]
    packed = [ops.code2x8_prepack(d["codes"]) for d in dev]
    ref = [ops.code2x8_matmat(d["x"], d["codes"], d["codebooks"], d["scales"], None) for d in dev]
    torch.cuda.synchronize()
    for rep in range(4):
        outs = [ops.code2x8_matmat_packed(d["x"], pk, O, d["codebooks"], d["scales"], None, x_ready=True)
                for d, pk in zip(dev, packed)]
        for o, r in zip(outs, ref):
            assert torch.equal(o, r)
    torch.cuda.synchronize()
    for ws in ops._workspaces.values():
        assert torch.count_nonzero(ws) == 0, "split-k accumulator words must be handed back zeroed"


def test_quantized_linear_module(pkg, orc):
    from rvv_lut_vq_b200 import QuantizedLinear, synth
    I, O = 512, 192
    L = synth.make_layer_np(I, O, seed=11, scales="randn", bias=True, batch=4)
    m = QuantizedLinear(I, O, 8, 1, 2, 8, bias=True, device="cuda", dtype=torch.float16)
    m.load_state_dict({"codes": cu(L["codes"]), "codebooks": cu(L["codebooks"]), "scales": cu(L["scales"]), "bias": cu(L["bias"])})
    y = m(cu(L["x"]))                                                 # batch 4 <= 6 -> gemv op
    assert y.shape == (4, O)
    for b in range(4):
        want = orc.c_aqlm_gemv_f64(L["x"][b], L["codes"], L["codebooks"]) * L["scales"].reshape(-1).astype(np.float64) + L["bias"]
        assert rel_err(y[b].float().cpu().numpy(), want) <= TOL_GEMV


@pytest.mark.parametrize("I,Os", [(4096, [4096, 4096, 4096]), (4096, [11008, 11008]), (8192, [8192, 1024, 1024]), (1024, [96, 1000]), (11008, [4096])])
def test_grouped_launch_is_bit_identical(pkg, ops, I, Os):
    """lutvq_code2x8_matmat_group_packed: n linears sharing x in one launch == n separate launches, bit for bit"""
    from rvv_lut_vq_b200 import synth
    Ls = [synth.make_layer_np(I, O, seed=100 + i, scales="randn", bias=(i % 2 == 0)) for i, O in enumerate(Os)]
    x = cu(Ls[0]["x"])
    packed = [ops.code2x8_prepack(cu(L["codes"])) for L in Ls]
    cbs = [cu(L["codebooks"]) for L in Ls]
    scs = [cu(L["scales"]) for L in Ls]
    bis = [None if L["bias"] is None else cu(L["bias"]) for L in Ls]
    want = [ops.code2x8_matmat_packed(x, packed[i], Os[i], cbs[i], scs[i], bis[i])[0] for i in range(len(Os))]
    for rep in range(3):                                                   # the workspace must come back all-zero
        got = ops.code2x8_matmat_group_packed(x, packed, Os, cbs, scs, bis)
        torch.cuda.synchronize()
        for i in range(len(Os)):
            assert torch.equal(got[i].view(torch.int16), want[i].view(torch.int16)), f"linear {i} of the group differs (rep {rep})"


# ------------------------------------------------------------------ Kx8 generalisation (N = 1, 2, 4 codebooks)
@pytest.mark.parametrize("I,O", [(11008, 4096), (1024, 96), (28672, 8192)])
def test_fused_swiglu_input_is_bit_identical(pkg, ops, I, O):
    """lutvq_code2x8_matmat_group_packed_swiglu == lutvq_decode_silu_mul followed by the plain grouped launch, bit for bit;
    the activation itself against torch (fp32 silu * up, one rounding)."""
    import ctypes
    from rvv_lut_vq_b200 import synth, _lib
    L = synth.make_layer_np(I, O, seed=I + O, scales="randn", bias=True)
    codes, cb, sc, bi = cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"]), cu(L["bias"])
    packed = ops.code2x8_prepack(codes)
    g = torch.Generator(device="cuda").manual_seed(5)
    gate = (torch.randn(I, device="cuda", generator=g) * 2).half()
    up = torch.randn(I, device="cuda", generator=g).half()
    act = torch.empty_like(up)
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    _lib.check(_lib.lib().lutvq_decode_silu_mul(P(gate), P(up), P(act), I, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    ref_act = (torch.nn.functional.silu(gate.float()) * up.float())
    assert (act.float() - ref_act).abs().max() <= 2e-3 * ref_act.abs().max()
    want = ops.code2x8_matmat_group_packed(act, [packed], [O], [cb], [sc], [bi])[0]
    got = ops.code2x8_matmat_group_packed(up, [packed], [O], [cb], [sc], [bi], gate=gate)[0]
    torch.cuda.synchronize()
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))


@pytest.mark.parametrize("name", ["aqlm_1x8_i128_o32", "aqlm_4x8_i128_o48", "aqlm_2x8_i256_o96"])
def test_codeKx8_vs_golden(pkg, ops, orc, golden_dir, name):
    """lutvq_codeKx8_matmat / _dequant against the fixtures generated from the reference's own numba_gemm_lut and
    _dequantize_weight (oracle/gen_golden.py), and bit-exact against the sequential-fp16-add restatement of CodeKx8Dequant."""
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    I, O, N, B, _, wb = [int(v) for v in g["cfg"]]
    x16 = g["x"].astype(np.float16)
    cb16 = g["codebooks"].astype(np.float16)
    sc16 = g["scales"].astype(np.float16)
    bias = g["bias"].astype(np.float16) if wb else None
    y = ops.codeKx8_matmat(cu(x16), cu(g["codes"]), cu(cb16), cu(sc16), None if bias is None else cu(bias))
    assert y.shape == (B, O)
    # float64 evaluation of the same fp16 operands: north-star tolerance 1e-3 of the layer's range
    Wd = np.zeros((O, I), dtype=np.float64)
    for n in range(N):
        Wd += cb16[n, g["idx"][:, :, n].astype(np.int64), 0, :].astype(np.float64).reshape(O, I)
    want = (x16.astype(np.float64) @ Wd.T) * sc16.reshape(-1).astype(np.float64) + (0 if bias is None else bias.astype(np.float64))
    got = y.float().cpu().numpy().astype(np.float64)
    assert np.abs(got - want).max() / np.abs(want).max() <= TOL_GEMV
    # the reference's fp32 numba result (fp32 operands) agrees to fp16-operand precision
    assert np.abs(got - g["y_numba"]).max() / np.abs(g["y_numba"]).max() <= 5e-3
    # dequantize: sequential fp16 adds in codebook order (cuda_kernel.cu:437-453), one more RN multiply for the scale
    W = ops.codeKx8_dequant(cu(g["codes"]), cu(cb16), None)
    seq = cb16[0, g["idx"][:, :, 0].astype(np.int64), 0, :]
    for n in range(1, N):
        seq = (seq + cb16[n, g["idx"][:, :, n].astype(np.int64), 0, :]).astype(np.float16)
    assert np.array_equal(bits(W), seq.reshape(O, I).view(np.uint16)), "CodeKx8Dequant semantics: not bit-exact"
    Ws = ops.codeKx8_dequant(cu(g["codes"]), cu(cb16), cu(sc16))
    seqs = (seq.reshape(O, I) * sc16.reshape(-1, 1)).astype(np.float16)
    assert np.array_equal(bits(Ws), seqs.view(np.uint16))
    ref = g["weight_f16_noscale"].astype(np.float16)
    if N <= 2:
        assert np.array_equal(bits(W), ref.view(np.uint16))              # one IEEE add: identical to _dequantize_weight
        assert torch.equal(W, ops.code2x8_dequant(cu(g["codes"]), cu(cb16), None)) if N == 2 else True
    else:
        np.testing.assert_allclose(W.float().cpu().numpy(), ref.astype(np.float32), rtol=2e-3, atol=2e-3)


# ------------------------------------------------------------------------------------------------ integer LUT (a13)
@pytest.mark.parametrize("cb,n_entries,n_lut,n_runs", [(16, 16, 1000, 37), (32, 16, 777, 5), (64, 64, 1000, 300), (16, 256, 10, 1)])
def test_simd_lut_lookup_add_vs_oracle(pkg, ops, orc, cb, n_entries, n_lut, n_runs):
    """bit-exact int16 wraparound sums against the scalar restatement (bench_avx_lut.cpp:98-107), non-zero start values"""
    rng = np.random.default_rng(cb * n_lut + n_runs)
    idx = rng.integers(0, n_entries, size=(n_lut, cb), dtype=np.uint8)
    lut = ((np.arange(n_entries) * 10 + 100) & 0xFF).astype(np.uint8).view(np.int8)     # the reference's table, wrapped to int8
    out0 = rng.integers(-30000, 30000, size=cb).astype(np.int16)
    base = orc.c_simd_lut_scalar(idx, lut, n_runs=n_runs)                               # starts from zero
    want = (out0.view(np.uint16).astype(np.uint32) + base.view(np.uint16).astype(np.uint32)).astype(np.uint16).view(np.int16)
    got = ops.simd_lut_lookup_add(cu(idx), cu(lut), cu(out0.copy()), n_runs)
    torch.cuda.synchronize()
    assert np.array_equal(got.cpu().numpy(), want)
    again = ops.simd_lut_lookup_add(cu(idx), cu(lut), cu(out0.copy()), n_runs)           # workspace came back zeroed
    assert np.array_equal(again.cpu().numpy(), want)


def test_simd_lut_full_size_is_n_runs_times_one_run(pkg, ops, orc):
    """the reference's own size (CB 16, 10 000 rows, 100 000 runs = 1.6e10 lookups): equals n_runs x (one oracle run) mod 2^16"""
    rng = np.random.default_rng(3)
    idx = rng.integers(0, 16, size=(10000, 16), dtype=np.uint8)
    lut = ((np.arange(16) * 10 + 100) & 0xFF).astype(np.uint8).view(np.int8)
    one = orc.c_simd_lut_scalar(idx, lut, n_runs=1).view(np.uint16).astype(np.uint64)
    want = ((one * 100000) & 0xFFFF).astype(np.uint16).view(np.int16)
    got = ops.simd_lut_lookup_add(cu(idx), cu(lut), torch.zeros(16, dtype=torch.int16, device="cuda"), 100000)
    torch.cuda.synchronize()
    assert np.array_equal(got.cpu().numpy(), want)


# ------------------------------------------------------------------ torch.ops.aqlm.* registration and module dispatch (a11)
def test_torch_ops_aqlm_registered_and_equal(pkg, ops):
    """the reference's op names (cuda_kernel.py:13-132) resolve to liblutvq and agree bit for bit with the python front end"""
    from rvv_lut_vq_b200 import synth, torch_ops  # noqa: F401
    L = synth.make_layer_np(512, 96, seed=3, scales="randn", bias=True, batch=3)
    x, codes, cb, sc, bi = cu(L["x"]), cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"]), cu(L["bias"])
    for name in ("code1x16_matmat", "code1x16_matmat_dequant", "code1x16_matmat_dequant_transposed", "code2x8_matmat", "code2x8_matmat_dequant",
                 "code2x8_matmat_dequant_transposed", "code1x8_matmat", "code1x8_matmat_dequant", "code1x8_matmat_dequant_transposed"):
        assert hasattr(torch.ops.aqlm, name), name
    y_op = torch.ops.aqlm.code2x8_matmat(x, codes, cb, sc, bi)
    assert torch.equal(y_op.view(torch.int16), ops.code2x8_matmat(x, codes, cb, sc, bi).view(torch.int16))
    y_nb = torch.ops.aqlm.code2x8_matmat(x, codes, cb, sc, None)                      # bias-free layers pass None
    assert torch.equal(y_nb.view(torch.int16), ops.code2x8_matmat(x, codes, cb, sc, None).view(torch.int16))
    xb = torch.randn((16, 512), dtype=torch.float16, device="cuda")
    assert torch.equal(torch.ops.aqlm.code2x8_matmat_dequant(xb, codes, cb, sc, bi).view(torch.int16),
                       ops.code2x8_matmat_dequant(xb, codes, cb, sc, bi).view(torch.int16))
    # transposed op = (g * scales) @ W   (cuda_kernel.cpp:486-519)
    g = torch.randn((2, 96), dtype=torch.float16, device="cuda")
    W = ops.code2x8_dequant(codes, cb, None)
    want = (g.double() * sc.reshape(1, -1).double()) @ W.double()
    got = torch.ops.aqlm.code2x8_matmat_dequant_transposed(g, codes, cb, sc, None)
    assert rel_err(got.double().cpu().numpy(), want.cpu().numpy()) <= 4e-3
    with pytest.raises((NotImplementedError, RuntimeError)):                            # no CPU implementation is registered
        torch.ops.aqlm.code2x8_matmat(x.cpu(), codes.cpu(), cb.cpu(), sc.cpu(), None)


@pytest.mark.parametrize("N,I,O,batch,dtype", [(1, 4096, 4096, 1, "f16"), (1, 11008, 4096, 2, "bf16"), (4, 4096, 4096, 1, "f16"), (4, 4096, 11008, 3, "bf16"),
                                               (8, 4096, 1024, 1, "f16"), (8, 1024, 515, 2, "bf16"), (4, 32, 33, 1, "f16"), (1, 128, 7, 1, "f16"),
                                               (3, 512, 96, 2, "f16"), (1, 192, 40, 1, "f16")])
def test_codeKx8_fused_vs_oracle(pkg, ops, orc, N, I, O, batch, dtype):
    """N in {1, 4, 8} codebooks through the fused LUT-GEMV (gemv_lut.cu templated on the codebooks per group), Llama-sized and
    ragged shapes, fp16 and bf16, against the float64 evaluation of the same operands (oracle C), scale + bias in the
    reference's rounding order.  N = 3 and I = 192 (24 code bytes per row) stay on the unfused path: same bar."""
    rng = np.random.default_rng(N * 100000 + I + O)
    dt = 0 if dtype == "f16" else 1
    tdt = torch.float16 if dt == 0 else torch.bfloat16
    idx = rng.integers(0, 256, size=(O, I // 8, N))
    codes = orc.np_pack_int_data(idx, 8)
    cb = torch.from_numpy(rng.standard_normal((N, 256, 1, 8)).astype(np.float32)).to(tdt)
    x = torch.from_numpy(rng.standard_normal((batch, I)).astype(np.float32)).to(tdt)
    sc = torch.from_numpy(rng.standard_normal((O, 1, 1, 1)).astype(np.float32)).to(tdt)
    bi = torch.from_numpy(rng.standard_normal(O).astype(np.float32)).to(tdt)
    tb = lambda t: t.contiguous().view(torch.int16).numpy().view(np.uint16)
    ones = torch.ones((O, 1, 1, 1), dtype=tdt)
    y = ops.codeKx8_matmat(x.cuda(), cu(codes), cb.cuda(), sc.cuda(), bi.cuda())
    y_acc = ops.codeKx8_matmat(x.cuda(), cu(codes), cb.cuda(), ones.cuda(), None)     # round(acc) * 1: the accumulator itself
    assert y.shape == (batch, O)
    for b in range(batch):
        acc = orc.c_aqlm_gemv_f64(tb(x[b]), codes, tb(cb), dt)
        assert rel_err(y_acc[b].double().cpu().numpy(), acc) <= (TOL_GEMV if dt == 0 else 8e-3)      # bf16 outputs carry 8 mantissa bits
        # the epilogue is round(acc) * scale + bias in the storage type (cuda_kernel.cpp:95-111): bit-exact from OUR rounded accumulator
        want = orc.c_scale_bias_epilogue(tb(y_acc[b].cpu()).reshape(1, -1), tb(sc.reshape(-1)), tb(bi), dt).reshape(-1)
        assert np.array_equal(tb(y[b].cpu()), want)
    if dt == 1 and N in (1, 4, 8):                                   # bf16 dequantize: sequential adds in the storage type
        W = ops.codeKx8_dequant(cu(codes), cb.cuda(), None)
        seq = cb[0, torch.from_numpy(idx[:, :, 0]), 0, :]
        for n in range(1, N):
            seq = seq + cb[n, torch.from_numpy(idx[:, :, n]), 0, :]
        assert torch.equal(W.cpu().view(torch.int16), seq.reshape(O, I).contiguous().view(torch.int16))


def test_codeKx8_fused_matches_2x8_kernel_bitwise(pkg, ops, orc):
    """N = 2 through lutvq_codeKx8_matmat is the very kernel of lutvq_code2x8_matmat: identical bits"""
    from rvv_lut_vq_b200 import synth
    L = synth.make_layer_np(1024, 320, seed=4, scales="randn", bias=True)
    a = ops.codeKx8_matmat(cu(L["x"]), cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"]), cu(L["bias"]))
    b = ops.code2x8_matmat(cu(L["x"]), cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"]), cu(L["bias"]))
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))


@pytest.mark.parametrize("fmt", ["1x8", "4x8", "1x16g8", "1x16g16"])
def test_quantized_linear_other_formats(pkg, ops, orc, fmt):
    """QuantizedLinear dispatches on (num_codebooks, nbits, group) like kernel_selector.py:21-102; gemv op and gemm op agree with float64"""
    from rvv_lut_vq_b200 import QuantizedLinear
    I, O = 512, 160
    N, nbits, gs = {"1x8": (1, 8, 8), "4x8": (4, 8, 8), "1x16g8": (1, 16, 8), "1x16g16": (1, 16, 16)}[fmt]
    rng = np.random.default_rng(9)
    idx = rng.integers(0, 2 ** nbits, size=(O, I // gs, N))
    m = QuantizedLinear(I, O, gs, 1, N, nbits, bias=True, device="cuda", dtype=torch.float16)
    cbk = torch.from_numpy(rng.standard_normal((N, 2 ** nbits, 1, gs)).astype(np.float16))
    sc = torch.from_numpy(rng.standard_normal((O, 1, 1, 1)).astype(np.float16))
    bi = torch.from_numpy(rng.standard_normal(O).astype(np.float16))
    m.load_state_dict({"codes": cu(orc.np_pack_int_data(idx, nbits)), "codebooks": cbk.cuda(), "scales": sc.cuda(), "bias": bi.cuda()})
    W64 = torch.zeros((O, I // gs, gs), dtype=torch.float64)
    for n in range(N):
        W64 += cbk[n, torch.from_numpy(idx[:, :, n]), 0, :].double()
    W64 = W64.reshape(O, I)
    for batch in (3, 12):                                                            # gemv op / gemm op (inference.py:95-96)
        x = torch.from_numpy(rng.standard_normal((batch, I)).astype(np.float16))
        y = m(x.cuda())
        want = (x.double() @ W64.T) * sc.reshape(1, -1).double() + bi.double()
        assert y.shape == (batch, O)
        assert rel_err(y.double().cpu().numpy(), want.numpy()) <= 3e-3                 # 1e-3 on the accumulator + the epilogue's fp16 roundings


def test_quantized_linear_repacks_after_reload_and_backward(pkg, ops):
    """the cached GEMV image follows the codes (reload / in-place edit), and grad_input flows through the transposed op (inference.py:99-142)"""
    from rvv_lut_vq_b200 import QuantizedLinear, synth
    I, O = 256, 64
    A, B = synth.make_layer_np(I, O, seed=1, scales="randn", bias=True), synth.make_layer_np(I, O, seed=2, scales="randn", bias=True)
    m = QuantizedLinear(I, O, 8, 1, 2, 8, bias=True, device="cuda", dtype=torch.float16)
    sd = lambda L: {"codes": cu(L["codes"]), "codebooks": cu(L["codebooks"]), "scales": cu(L["scales"]), "bias": cu(L["bias"])}
    x = cu(A["x"])
    m.load_state_dict(sd(A))
    ya = m(x)
    m.load_state_dict(sd(B))                                                         # same storage, new contents: _version changes
    yb = m(x)
    assert torch.equal(yb.view(torch.int16), ops.code2x8_matmat(x, cu(B["codes"]), cu(B["codebooks"]), cu(B["scales"]), cu(B["bias"])).view(torch.int16))
    assert not torch.equal(ya, yb)
    xg = x.clone().requires_grad_(True)
    with torch.enable_grad():
        y = m(xg)
        y.float().sum().backward()
    W = ops.code2x8_dequant(cu(B["codes"]), cu(B["codebooks"]), cu(B["scales"]))
    want = W.double().sum(dim=0)
    assert rel_err(xg.grad.double().cpu().numpy().reshape(-1), want.cpu().numpy()) <= 5e-3


def test_prefill_chain_respects_stream_order(pkg, ops):
    """The prefill kernel is launched with programmatic stream serialization and starts on its weights before the previous kernel
    has finished (gemm_tc.cu: everything before griddepcontrol.wait).  A chain in which every call's input is the previous call's
    output, issued back to back, must give the bits of the same chain with a device synchronisation after every call."""
    from rvv_lut_vq_b200 import synth
    dev = torch.device("cuda", 0)
    for (n, batch, dtype) in [(1024, 16, torch.float16), (2048, 200, torch.float16), (1024, 64, torch.bfloat16)]:
        Ls = [synth.make_layer_torch(n, n, dev, dtype=dtype, seed=90 + k, scales="ones") for k in range(3)]
        x = torch.randn((batch, n), device=dev, dtype=torch.float32).to(dtype)
        for L in Ls:                                     # scale every layer so that the chain stays O(1): weights ~ N(0, 2), n terms
            L["scales"] = (L["scales"].float() / (2.0 * n) ** 0.5).to(dtype)

        def chain(sync):
            y = x
            for rep in range(4):
                for L in Ls:
                    y = ops.code2x8_matmat_dequant_fused(y, L["codes"], L["codebooks"], L["scales"], None)
                    if sync:
                        torch.cuda.synchronize()
            torch.cuda.synchronize()
            return y
        y_sync = chain(True)
        assert torch.isfinite(y_sync.float()).all() and float(y_sync.float().abs().max()) > 1e-3
        for _ in range(3):
            assert torch.equal(chain(False).view(torch.int16), y_sync.view(torch.int16))
        # eager launches are ~20 us apart; inside a CUDA graph the 12 launches really run back to back (the bench's situation)
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            chain(False)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                y = x
                for rep in range(4):
                    for L in Ls:
                        y = ops.code2x8_matmat_dequant_fused(y, L["codes"], L["codebooks"], L["scales"], None)
            for _ in range(5):
                g.replay()
                st.synchronize()
                assert torch.equal(y.view(torch.int16), y_sync.view(torch.int16))
        torch.cuda.synchronize()
