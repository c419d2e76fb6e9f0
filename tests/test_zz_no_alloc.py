"""lutvq_code2x8_matmat_noalloc (csrc/noalloc.cu): the fused LUT-GEMV with its split-k words in a static __device__ array inside
the library -- no workspace argument, no allocation in the call.  Called through the C-ABI with raw pointers and compared with
the existing entry point lutvq_code2x8_matmat: same kernel, integer split-k sums => the bits must be identical (and, as a looser
second bar, within the 1e-3 tolerance the GEMV tests use against the float64 oracle).  Two calls in a row on different shapes,
so the second one finds the array as the first left it.  Named test_zz_* so pytest collects it last."""
import ctypes
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL_GEMV = 1e-3


@pytest.fixture(scope="module")
def abi():
    import rvv_lut_vq_b200 as r
    assert os.path.exists(r._lib.LIB_PATH), "liblutvq.so must be prebuilt in-tree (no fallback)"
    assert torch.cuda.is_available()
    return r._lib


def P(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def run_pair(abi, I, O, batch, seed, dtype="f16", bias=True):
    """-> (y from the no-allocation entry point, y from lutvq_code2x8_matmat, layer)"""
    from rvv_lut_vq_b200 import synth
    lib = abi.lib()
    L = synth.make_layer_np(I, O, seed=seed, scales="randn", bias=bias, batch=batch)
    conv = cu if dtype == "f16" else (lambda a: cu(a.astype(np.float32)).to(torch.bfloat16))
    dt = abi.LUTVQ_F16 if dtype == "f16" else abi.LUTVQ_BF16
    x, cb, sc = conv(L["x"]), conv(L["codebooks"]), conv(L["scales"]).reshape(-1).contiguous()
    bi = conv(L["bias"]) if bias else None
    codes = cu(L["codes"])
    y_new = torch.full((batch, O), float("nan"), dtype=x.dtype, device="cuda")
    y_old = torch.full((batch, O), float("nan"), dtype=x.dtype, device="cuda")
    ws = torch.zeros(int(lib.lutvq_workspace_bytes(O, I)), dtype=torch.uint8, device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    abi.check(lib.lutvq_code2x8_matmat_noalloc(P(x), P(codes), P(cb), P(sc), P(bi), P(y_new), batch, O, I, dt, stream))
    abi.check(lib.lutvq_code2x8_matmat(P(x), P(codes), P(cb), P(sc), P(bi), P(y_old), batch, O, I, dt, P(ws), ws.numel(), stream))
    torch.cuda.synchronize()
    return y_new, y_old, L


def test_noalloc_matches_matmat_two_calls_in_a_row(abi, orc):
    # call 1: Llama-2-7B gate/up shape (16 k-slices, 344 row sets); call 2: another shape on the array call 1 used
    for (I, O, batch, seed, dtype) in [(4096, 11008, 2, 11, "f16"), (1024, 515, 3, 12, "f16"), (11008, 4096, 1, 13, "bf16")]:
        y_new, y_old, L = run_pair(abi, I, O, batch, seed, dtype)
        assert not torch.isnan(y_new.float()).any()
        assert torch.equal(y_new.view(torch.int16), y_old.view(torch.int16)), f"{I}x{O}: differs from lutvq_code2x8_matmat"
        if dtype == "f16":          # and the float64 oracle, at the tolerance of tests/test_gpu_parity.py
            for b in range(batch):
                y64 = orc.c_aqlm_gemv_f64(L["x"][b], L["codes"], L["codebooks"]) * L["scales"].reshape(-1).astype(np.float64) + L["bias"]
                got = y_new[b].double().cpu().numpy()
                assert np.abs(got - y64).max() / np.abs(y64).max() <= 2 * TOL_GEMV      # + the fp16 roundings of the epilogue
    # the same call twice: bit-identical to itself (the array is cleared per call, nothing carries over)
    a, _, _ = run_pair(abi, 4096, 4096, 1, 21)
    b, _, _ = run_pair(abi, 4096, 4096, 1, 21)
    assert torch.equal(a.view(torch.int16), b.view(torch.int16))


def test_noalloc_limit_and_errors(abi):
    lib = abi.lib()
    max_rows = 32768                                                   # LUTVQ_NOALLOC_MAX_ROWS (include/lutvq.h)
    assert int(lib.lutvq_noalloc_scratch_bytes()) == 256 + 8 * max_rows
    z = torch.zeros(64, dtype=torch.float16, device="cuda")
    codes = torch.zeros(16, dtype=torch.int8, device="cuda")
    cb = torch.zeros((2, 256, 1, 8), dtype=torch.float16, device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.lutvq_code2x8_matmat_noalloc(P(z), P(codes), P(cb), None, None, P(z), 1, max_rows + 1, 64, 0, stream)
    assert rc == abi.LUTVQ_EINVAL
    msg = lib.lutvq_last_error().decode()
    assert "LUTVQ_NOALLOC_MAX_ROWS" in msg and str(max_rows) in msg
    assert lib.lutvq_code2x8_matmat_noalloc(P(z), P(codes), P(cb), None, None, P(z), 1, 8, 40, 0, stream) == abi.LUTVQ_EINVAL    # prob_k % 64
    assert lib.lutvq_code2x8_matmat_noalloc(P(z), P(codes), P(cb), None, None, P(z), 1, 8, 64, 7, stream) == abi.LUTVQ_EUNSUPPORTED
    assert lib.lutvq_code2x8_matmat_noalloc(None, P(codes), P(cb), None, None, None, 0, 8, 64, 0, stream) == 0                  # empty batch


def test_noalloc_at_the_row_limit(abi):
    """prob_m = LUTVQ_NOALLOC_MAX_ROWS exactly fills the array"""
    y_new, y_old, _ = run_pair(abi, 256, 32768, 1, 31, bias=False)
    assert torch.equal(y_new.view(torch.int16), y_old.view(torch.int16))
