#!/usr/bin/env python
"""Sweep the GEMV kernel's launch knobs (warps/CTA, prefetch depth, CTAs/SM, PDL) over the Llama-2-7B shapes.
Each measurement replays a CUDA graph of GEMVs over a ring of distinct layers (> 2x L2), CUDA-event timed.
    python tools/tune_gemv.py [--shapes 4096x11008,...] > gpurun_out/tune.jsonl"""
import argparse
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rvv_lut_vq_b200 as pkg
from rvv_lut_vq_b200 import ops, synth, _lib


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="4096x11008,4096x4096,11008x4096")
    ap.add_argument("--tunings", default="8,2,1;8,2,2;8,3,1;8,1,2;16,1,1;16,2,1;4,2,2;4,4,2;4,2,3;8,3,2")
    ap.add_argument("--reps", type=int, default=30)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    lib = pkg.lib()
    stream = torch.cuda.Stream(dev)
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    peak = 6539.9
    for shape in args.shapes.split(","):
        i_f, o_f = [int(v) for v in shape.split("x")]
        nb = max(8, int(280e6 // (o_f * i_f // 4)))
        Ls = [synth.make_layer_torch(i_f, o_f, dev, seed=k, scales="ones") for k in range(nb)]
        xv = torch.randn((nb, 1, i_f), device=dev, dtype=torch.float16)
        yv = torch.empty((nb, 1, o_f), device=dev, dtype=torch.float16)
        nbytes = synth.gemv_algorithmic_bytes(i_f, o_f)
        for tun in args.tunings.split(";"):
            nw, pf, cps = [int(v) for v in tun.split(",")]
            lib.lutvq_debug_set_gemv_tuning(nw, pf, cps)
            for flags in (0, _lib.FLAG_NO_PDL):
                with torch.cuda.stream(stream):
                    ws = ops.workspace(dev, o_f, i_f)
                    ws.zero_()

                    def body():
                        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
                        for k, L in enumerate(Ls):
                            _lib.check(lib.lutvq_code2x8_matmat_ex(P(xv[k]), P(L["codes"]), P(L["codebooks"]), P(L["scales"]), None,
                                                                   P(yv[k]), 1, o_f, i_f, 0, flags, P(ws), ws.numel(), st))
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
                    for _ in range(args.reps):
                        g.replay()
                    b.record(stream)
                    torch.cuda.synchronize()
                    us = a.elapsed_time(b) * 1e3 / (args.reps * nb)
                    nk, nc, rpc = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
                    lib.lutvq_debug_get_gemv_plan(o_f, i_f, ctypes.byref(nk), ctypes.byref(nc), ctypes.byref(rpc))
                    print(json.dumps({"shape": shape, "nw": nw, "pf": pf, "ctas_per_sm": cps, "pdl": flags == 0, "us": round(us, 3),
                                      "gbps": round(nbytes / us / 1e3, 1), "frac": round(nbytes / us / 1e3 / peak, 4),
                                      "grid": [nc.value, nk.value], "rows_per_chunk": rpc.value}), flush=True)
                    del g
        del Ls


if __name__ == "__main__":
    main()
