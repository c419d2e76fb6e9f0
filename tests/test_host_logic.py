"""CPU: host-side logic -- layout proof, C-ABI surface, loud failure without a GPU, reference-mirroring API."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def pkg():
    import rvv_lut_vq_b200 as r
    r.build()
    return r


def test_layout_proof(tmp_path):
    """tests/layout_check.cpp: PRMT addressing, bank-conflict freedom, build<->lookup agreement, warp reduction."""
    exe = tmp_path / "layout_check"
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    subprocess.check_call([cxx, "-O1", "-std=c++17", "-o", str(exe), os.path.join(ROOT, "tests", "layout_check.cpp")])
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0 and "layout OK" in out.stdout, out.stdout + out.stderr


def test_abi_exports_every_declared_symbol(pkg):
    hdr = open(os.path.join(ROOT, "include", "lutvq.h")).read()
    declared = set(re.findall(r"\b(lutvq_[A-Za-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    lib = ctypes.CDLL(pkg._lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"liblutvq.so does not export {name}"
    assert declared == set(pkg._lib.SIGNATURES), "python signature table out of sync with include/lutvq.h"


def test_abi_is_plain_c(pkg):
    """exported lutvq_* symbols are unmangled and the library does not link torch"""
    nm = subprocess.run(["nm", "-D", "--defined-only", pkg._lib.LIB_PATH], capture_output=True, text=True).stdout
    syms = [l.split()[-1] for l in nm.splitlines() if " T " in l]
    assert "lutvq_code2x8_matvec" in syms
    ldd = subprocess.run(["ldd", pkg._lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in ldd and "c10" not in ldd


def test_version_and_workspace(pkg):
    lib = pkg.lib()
    assert lib.lutvq_version() >= 100
    # split-k words (8 bytes per padded output row) + a 256-byte reserved header, for packed and reference-layout codes alike
    # (the kernel reads either layout directly: no staging area)
    for O, I in [(4096, 4096), (11008, 4096), (4096, 11008), (96, 256)]:
        small = lib.lutvq_packed_workspace_bytes(O, I)
        O_pad = -(-O // 32) * 32
        assert small == 256 + O_pad * 8
        need = lib.lutvq_workspace_bytes_ex(O, I, 6)                     # decode-only scratch
        assert need == small
        assert lib.lutvq_workspace_bytes(O, I) >= need                 # any-batch scratch also covers the prefill split-k partials
        assert lib.lutvq_code2x8_packed_bytes(O, I) == -(-I // 256) * (O_pad // 32) * 2048
    assert lib.lutvq_workspace_bytes(0, 64) == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu(pkg):
    lib = pkg.lib()
    buf = np.zeros(1 << 16, dtype=np.uint8)
    p = ctypes.c_void_p(buf.ctypes.data - buf.ctypes.data % 16 + 16)
    rc = lib.lutvq_code2x8_matvec(p, p, p, p, 64, 64, 0, p, 1 << 15, None)
    assert rc == pkg._lib.LUTVQ_ECUDA
    assert b"no CPU fallback" in lib.lutvq_last_error()
    from rvv_lut_vq_b200 import ops
    x = torch.zeros(1, 64, dtype=torch.float16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.code2x8_matmat(x, torch.zeros(8, 8, 2, dtype=torch.int8), torch.zeros(2, 256, 1, 8, dtype=torch.float16),
                           torch.ones(8, 1, 1, 1, dtype=torch.float16), None)


def test_argument_validation(pkg):
    lib = pkg.lib()
    buf = np.zeros(1 << 12, dtype=np.uint8)
    a = buf.ctypes.data - buf.ctypes.data % 16 + 16
    p = ctypes.c_void_p(a)
    E = pkg._lib.LUTVQ_EINVAL
    assert lib.lutvq_code2x8_matvec(None, p, p, p, 64, 64, 0, p, 64, None) == E
    assert lib.lutvq_code2x8_matvec(p, p, p, p, 64, 72, 0, p, 64, None) == E           # prob_k % 64
    assert b"multiple of 64" in lib.lutvq_last_error()
    assert lib.lutvq_code2x8_matvec(ctypes.c_void_p(a + 4), p, p, p, 64, 64, 0, p, 64, None) == E   # alignment
    assert lib.lutvq_code2x8_matvec(p, p, p, p, 64, 64, 7, p, 64, None) == pkg._lib.LUTVQ_EUNSUPPORTED
    assert b"float16" in lib.lutvq_last_error()
    assert lib.lutvq_code2x8_dequant(p, None, p, None, 64, 64, 0, None) == E
    assert lib.lutvq_lut_build(p, p, p, 4, 2, 256, 64, 0, 0, None) == E                # d too large
    assert lib.lutvq_lut_gemv(p, p, p, 4, 2, 2, 256, 0, 9, None) == E                  # bad acc mode


def test_reference_mirroring_api():
    """names / signatures / parameter layout a user of the reference expects (inference.py:11-61, utils.py:11-31)"""
    from rvv_lut_vq_b200 import QuantizedLinear, ops, utils
    for name in ("code2x8_matmat", "code2x8_matvec", "code2x8_dequant", "code2x8_matmat_dequant"):
        assert callable(getattr(ops, name))
    m = QuantizedLinear(256, 96, in_group_size=8, out_group_size=1, num_codebooks=2, nbits_per_codebook=8, bias=True,
                        dtype=torch.float16)
    sd = {k: (tuple(v.shape), v.dtype) for k, v in m.state_dict().items()}
    assert sd == {
        "codebooks": ((2, 256, 1, 8), torch.float16),
        "codes": ((96, 32, 2), torch.int8),
        "scales": ((96, 1, 1, 1), torch.float16),
        "bias": ((96,), torch.float16),
    }
    assert m.use_gemv_rule(torch.zeros(1, 6, 256)) and not m.use_gemv_rule(torch.zeros(1, 7, 256))
    with pytest.raises(NotImplementedError):
        m(torch.zeros(1, 256, dtype=torch.float16))          # CPU tensor: no CPU path
    m16 = QuantizedLinear(256, 96, 8, 1, 1, 16, bias=False, dtype=torch.float16, device="meta")
    assert m16.codes.dtype == torch.int16 and m16.codebook_size == 65536
    # pack/unpack round trip incl. the reference's in-place wrap
    idx = torch.randint(256, (5, 7, 2))
    keep = idx.clone()
    packed = utils.pack_int_data(idx, 8)
    assert packed.dtype == torch.int8 and torch.equal(utils.unpack_int_data(packed, 8), keep)
    with pytest.raises(NotImplementedError):
        ops.check_use_bfloat16(torch.zeros(1, dtype=torch.float32))


def test_algorithmic_bytes_table():
    """SURVEY 8 table / BASELINE.md section 2"""
    from rvv_lut_vq_b200 import synth
    assert synth.gemv_algorithmic_bytes(4096, 4096) == 4227072
    assert synth.gemv_algorithmic_bytes(4096, 11008) == 11332608
    assert synth.gemv_algorithmic_bytes(11008, 4096) == 11318784


def test_traffic_counter_matches_reference_rows():
    """reference_dataflow_traffic replays VQ.dataflow_sim's loop nests (vq.py:68-267) on sizes only; every
    counter of every (loop order, configuration) row equals what the reference's own code produced (oracle/gen_golden_dataflow.py)."""
    import csv
    from rvv_lut_vq_b200 import traffic
    rows = list(csv.DictReader(open(os.path.join(ROOT, "tests", "golden", "dataflow_sim_ref.csv"))))
    assert len(rows) == 28
    for r in rows:
        cfg = (int(r["D"]), int(r["M"]), int(r["N"]), int(r["K"]))
        got = traffic.reference_dataflow_traffic(cfg, int(r["Din"]), int(r["Dout"]), r["Dataflow"], int(r["VLEN"]))
        for k in ("codeword", "codebook", "lut", "inp", "psum"):
            assert got[k] == float(r[k + "_MB"]), (r["VQ_type"], r["Dataflow"], k)
        assert (got["N_tile"], got["M_tile"], got["CW_tile"]) == (int(r["N_tile"]), int(r["M_tile"]), int(r["CW_tile"]))
    with pytest.raises(NotImplementedError):
        traffic.reference_dataflow_traffic((128, 8, 1, 256), 1, 1024, "OMND-compact")


def test_gpu_traffic_model_vs_measured_dram_bytes():
    """the model's HBM figure is DESIGN.md's algorithmic byte count, and the DRAM bytes ncu measured for the kernel
    (profiles/traffic.json) are within 2 % of it: no wasted re-reads"""
    import json
    from rvv_lut_vq_b200 import synth, traffic
    for (i, o) in [(4096, 4096), (4096, 11008), (11008, 4096)]:
        m = traffic.gemv_traffic_model(i, o)
        assert m["hbm"] == synth.gemv_algorithmic_bytes(i, o)
        assert m["smem_ld"] == 4 * m["hbm"] - 4 * (8192 + 2 * i + 4 * o) and m["n_k"] == -(-i // 256)
    measured = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["lut_gemv_2x8_kernel_4096x11008_dram_bytes_per_launch"]
    assert 1.0 <= measured / traffic.gemv_traffic_model(4096, 11008)["hbm"] <= 1.02


def test_hf_checkpoint_layout_round_trip(tmp_path):
    """the HF layout convert_to_hf.py:50-100 writes -- state-dict key names, packed code dtypes, config.json's
    quantization_config -- read back without transformers; malformed checkpoints fail loudly"""
    from rvv_lut_vq_b200 import hf_loader, utils
    cfg = {"model_type": "llama", "hidden_size": 64, "intermediate_size": 128, "num_hidden_layers": 1, "num_attention_heads": 2, "vocab_size": 50,
           "quantization_config": {"quant_method": "aqlm", "nbits_per_codebook": 8, "num_codebooks": 2, "out_group_size": 1, "in_group_size": 8,
                                   "linear_weights_not_to_quantize": ["model.embed_tokens.weight", "lm_head.weight"]}}
    g = torch.Generator().manual_seed(0)
    state = {}
    for name, (i, o) in {"self_attn.q_proj": (64, 64), "mlp.down_proj": (128, 64)}.items():
        p = "model.layers.0." + name
        state[p + ".codes"] = utils.pack_int_data(torch.randint(256, (o, i // 8, 2), generator=g), 8)   # convert_to_hf.py:66
        state[p + ".codebooks"] = torch.randn((2, 256, 1, 8), generator=g).half()
        state[p + ".scales"] = torch.randn((o, 1, 1, 1), generator=g).half()
    state["model.embed_tokens.weight"] = torch.randn((50, 64), generator=g).half()
    hf_loader.save_checkpoint(str(tmp_path), cfg, state)
    cfg2, state2 = hf_loader.load_checkpoint(str(tmp_path))
    assert cfg2["quantization_config"]["num_codebooks"] == 2
    assert set(state2) == set(state) and all(torch.equal(state2[k], state[k]) and state2[k].dtype == state[k].dtype for k in state)
    assert hf_loader.quantized_prefixes(state2) == ["model.layers.0.mlp.down_proj", "model.layers.0.self_attn.q_proj"]
    assert state2["model.layers.0.self_attn.q_proj.codes"].dtype == torch.int8
    dc = hf_loader.llama_decoder_config(cfg2)
    assert (dc.hidden, dc.intermediate, dc.n_layers, dc.n_heads, dc.n_kv_heads, dc.head_dim) == (64, 128, 1, 2, 2, 32)
    bad = dict(state)
    del bad["model.layers.0.self_attn.q_proj.scales"]
    with pytest.raises(KeyError):
        hf_loader.quantized_prefixes(bad)
    cfg_bad = dict(cfg)
    cfg_bad.pop("quantization_config")
    hf_loader.save_checkpoint(str(tmp_path / "noq"), cfg_bad, state)
    with pytest.raises(ValueError):
        hf_loader.load_checkpoint(str(tmp_path / "noq"))


def test_prefill_route_thresholds():
    """ops.two_step_min_rows: the fused tcgen05 kernel while the rows fit one launch (<= 256), the dequantize-once + library GEMM route above;
    small layers (<= 4096 x 4096 weights: the fused kernel splits k) switch from 129 rows (profiles/r02_prefill_shapes.jsonl)."""
    from rvv_lut_vq_b200 import ops
    assert ops.two_step_min_rows(11008, 4096) == ops.TWO_STEP_MIN_ROWS == 257
    assert ops.two_step_min_rows(4096, 11008) == 257 and ops.two_step_min_rows(12288, 4096) == 257
    assert ops.two_step_min_rows(4096, 4096) == ops.TWO_STEP_MIN_ROWS_SMALL == 129
    assert ops.two_step_min_rows(1024, 1024) == 129
    assert ops.TWO_STEP_MIN_ROWS_SMALL <= ops.TWO_STEP_MIN_ROWS
