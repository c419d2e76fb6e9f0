#!/usr/bin/env python
"""Time the reference's OWN CUDA kernels (unmodified, vendored + built by tools/vendor_ref_cuda.py into the git-ignored
baseline/_ref/aqlm_cuda/) on the same B200 beside ours: the "kernel to beat" of SURVEY 8d / BASELINE.md C5.

    python tools/ref_cuda_bench.py [--out gpurun_out/kernel_to_beat.json] [--quick]

Protocol = benchmark/kernel/rq_lut/benchmark_rq_gemm_gpu.py:79-116 (same tensors: x randn fp16, codes U{0..255} int8
[O, I/8, 2], codebooks randn [2,256,1,8], scales ones) but L2-cold and launch-overhead-free like bench.py: a ring of distinct
layers > 2x L2 per shape, one CUDA graph per ring, CUDA events.  Also an independent parity witness: the two
implementations' outputs are compared with each other and with a torch float64 dense evaluation of the dequantized weights.
Nothing here touches oracle/ or /root/reference."""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF_SO = os.path.join(ROOT, "baseline", "_ref", "aqlm_cuda", "build", "codebook_cuda.so")


def load_reference():
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    if not os.path.exists(REF_SO):
        return None, f"{REF_SO} missing: run tools/vendor_ref_cuda.py in the build container"
    try:
        spec = importlib.util.spec_from_file_location("codebook_cuda", REF_SO)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod, None
    except Exception as e:  # noqa: BLE001
        return None, f"import failed: {e!r}"


def time_graph(torch, stream, body, reps):
    body()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        body()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(reps):
        g.replay()
    b.record(stream)
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps            # us per graph replay


def run(quick: bool = False, shapes=None, batches=(16, 64, 256)):
    import torch
    import rvv_lut_vq_b200 as pkg
    from rvv_lut_vq_b200 import ops, synth

    ref, why = load_reference()
    if ref is None:
        return {"unavailable": why}
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = torch.cuda.Stream(dev)
    out = {"source": "baseline/_ref/aqlm_cuda (cuda_kernel.cu/.cpp of the reference, unmodified, -gencode arch=compute_100a,code=sm_100a)",
           "protocol": "ring of distinct layers > 2x L2 per shape, one CUDA graph per ring, CUDA events; batch-1 rows are us per GEMV",
           "gemv": [], "dequant": None, "prefill": []}
    shapes = shapes or synth.SWEEP_7B
    reps = 5 if quick else 20
    with torch.cuda.stream(stream):
        for (i_f, o_f) in shapes:
            nb = max(8, int(280e6 // (o_f * i_f // 4)) + 1)
            nb = -(-nb // 8) * 8
            if quick:
                nb = 8
            Ls = [synth.make_layer_torch(i_f, o_f, dev, seed=300 + k, scales="ones") for k in range(nb)]
            xs = torch.randn((nb, 1, 1, i_f), device=dev, dtype=torch.float16)
            packed = [ops.code2x8_prepack(L["codes"]) for L in Ls]
            res = {}

            def ref_body():
                for k, L in enumerate(Ls):
                    res[k] = ref.code2x8_matmat(xs[k], L["codes"], L["codebooks"], L["scales"], None)
            us_ref = time_graph(torch, stream, ref_body, reps) / nb
            y_ref = res[0].clone()

            def our_dropin():                              # same call shape, reference code layout (no prepack)
                for k, L in enumerate(Ls):
                    res[k] = ops.code2x8_matmat(xs[k], L["codes"], L["codebooks"], L["scales"], None)
            us_dropin = time_graph(torch, stream, our_dropin, reps) / nb
            y_our = res[0].clone()

            def our_packed():                              # QuantizedLinear's path: load-time prepack, dependent chain
                for k, L in enumerate(Ls):
                    res[k] = ops.code2x8_matmat_packed(xs[k], packed[k], o_f, L["codebooks"], L["scales"], None)
            us_packed = time_graph(torch, stream, our_packed, reps) / nb
            y_pk = res[0].clone()

            def our_xready():
                for k, L in enumerate(Ls):
                    res[k] = ops.code2x8_matmat_packed(xs[k], packed[k], o_f, L["codebooks"], L["scales"], None, x_ready=True)
            us_xr = time_graph(torch, stream, our_xready, reps) / nb
            # parity witness: both against float64 dense on the bit-exact dequantized weights
            W = ops.code2x8_dequant(Ls[0]["codes"], Ls[0]["codebooks"], None).double()
            y64 = (W @ xs[0].reshape(-1).double())
            den = float(y64.abs().max())
            row = {"shape": f"{i_f}->{o_f}", "layers_in_ring": nb, "reference_us": round(us_ref, 3),
                   "ours_dropin_us": round(us_dropin, 3), "ours_packed_dependent_us": round(us_packed, 3), "ours_packed_xready_us": round(us_xr, 3),
                   "speedup_dropin": round(us_ref / us_dropin, 2), "speedup_packed_dependent": round(us_ref / us_packed, 2),
                   "speedup_packed_xready": round(us_ref / us_xr, 2),
                   "max_rel_err_vs_f64": {"reference": float((y_ref.reshape(-1).double() - y64).abs().max()) / den,
                                          "ours": float((y_our.reshape(-1).double() - y64).abs().max()) / den,
                                          "ours_packed": float((y_pk.reshape(-1).double() - y64).abs().max()) / den},
                   "max_abs_diff_ours_vs_reference": float((y_our.float() - y_ref.float()).abs().max())}
            out["gemv"].append(row)
            print(json.dumps(row), flush=True)
            del Ls, packed, res

        # ---- dequantize + prefill on the headline shape ----
        i_f, o_f = 4096, 11008
        nb = 4 if quick else 8
        Ls = [synth.make_layer_torch(i_f, o_f, dev, seed=500 + k, scales="ones") for k in range(nb)]
        res = {}

        def ref_dq():
            for k, L in enumerate(Ls):
                res[k] = ref.code2x8_dequant(L["codes"], L["codebooks"], L["scales"])
        us = time_graph(torch, stream, ref_dq, reps) / nb
        w_ref = res[0].clone()

        def our_dq():
            for k, L in enumerate(Ls):
                res[k] = ops.code2x8_dequant(L["codes"], L["codebooks"], L["scales"])
        us_o = time_graph(torch, stream, our_dq, reps) / nb
        out["dequant"] = {"shape": f"{i_f}x{o_f}", "reference_us": round(us, 2), "ours_us": round(us_o, 2), "speedup": round(us / us_o, 2),
                          "bit_identical": bool(torch.equal(res[0].view(torch.int16), w_ref.view(torch.int16)))}
        print(json.dumps(out["dequant"]), flush=True)
        for b in batches:
            xb = torch.randn((nb, b, i_f), device=dev, dtype=torch.float16)

            def ref_pf():
                for k, L in enumerate(Ls):
                    res[k] = ref.code2x8_matmat_dequant(xb[k], L["codes"], L["codebooks"], L["scales"], None)
            us = time_graph(torch, stream, ref_pf, reps) / nb
            y_ref = res[0].clone()

            def our_pf():
                for k, L in enumerate(Ls):
                    res[k] = ops.code2x8_matmat_dequant(xb[k], L["codes"], L["codebooks"], L["scales"], None)
            us_o = time_graph(torch, stream, our_pf, reps) / nb
            den = float(y_ref.float().abs().max())
            row = {"batch": b, "reference_us": round(us, 2), "ours_us": round(us_o, 2), "speedup": round(us / us_o, 2),
                   "max_abs_diff_over_max": float((res[0].float() - y_ref.float()).abs().max()) / den}
            out["prefill"].append(row)
            print(json.dumps(row), flush=True)
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "kernel_to_beat.json"))
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    r = run(a.quick)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(r, f, indent=1)
    print(json.dumps(r))
