#!/usr/bin/env python
"""BASELINE.json configs[2]: codebook dequantize + tcgen05 GEMM at prefill batch sizes on 4096x11008 AQLM 2x8, and the
stand-alone dequantize kernel.  CUDA-event timed, ring of distinct layers (> 2x L2), JSON lines on stdout."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rvv_lut_vq_b200 import ops, synth  # noqa: E402


def timed(fn, reps):
    """reps calls captured in ONE CUDA graph (the Python wrapper costs more host time per call than the kernel runs)"""
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for _ in range(3):
            fn()
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for _ in range(reps):
                fn()
        g.replay()
        st.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(3):
            g.replay()
        b.record(st)
        st.synchronize()
    return a.elapsed_time(b) * 1e3 / (3 * reps)


def main():
    dev = torch.device("cuda", 0)
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    I, O, NB = 4096, 11008, 24
    Ls = [synth.make_layer_torch(I, O, dev, seed=k, scales="ones") for k in range(NB)]
    for batch in (16, 32, 64, 128, 256, 512, 1024, 2048, 4096):
        x = torch.randn((batch, I), device=dev, dtype=torch.float16)
        k = [0]

        def f():
            L = Ls[k[0] % NB]
            k[0] += 1
            ops.code2x8_matmat_dequant(x, L["codes"], L["codebooks"], L["scales"], None)
        us = timed(f, 48 if batch <= 256 else 12)
        flops = 2.0 * batch * I * O
        xw = x.float()
        W = ops.code2x8_dequant(Ls[0]["codes"], Ls[0]["codebooks"], None)
        us_cublas = timed(lambda: torch.nn.functional.linear(x, W), 48 if batch <= 256 else 12)
        us_deq = timed(lambda: ops.code2x8_dequant(Ls[k[0] % NB]["codes"], Ls[k[0] % NB]["codebooks"], None), 24)
        print(json.dumps({"op": "code2x8_matmat_dequant", "shape": f"{I}x{O}", "batch": batch, "us": round(us, 2),
                          "tflops": round(flops / us / 1e6, 1), "frac_of_bf16_burst_peak": round(flops / us / 1e6 / peaks["bf16_tflops"], 4),
                          "code_stream_GBps": round(O * I / 4 / us / 1e3, 1),
                          "reference_style_dequant_plus_cublas_us": round(us_deq + us_cublas, 2), "cublas_only_us": round(us_cublas, 2)}), flush=True)
    us = timed(lambda: ops.code2x8_dequant(Ls[0]["codes"], Ls[0]["codebooks"], Ls[0]["scales"]), 48)
    nbytes = O * I // 4 + 2 * O * I + 8192 + 2 * O
    print(json.dumps({"op": "code2x8_dequant", "shape": f"{I}x{O}", "us": round(us, 2), "GBps": round(nbytes / us / 1e3, 1),
                      "frac_of_hbm_peak": round(nbytes / us / 1e3 / peaks["hbm_gbs"], 4), "algorithmic_bytes": nbytes}))


if __name__ == "__main__":
    main()
