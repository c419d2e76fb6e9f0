#!/usr/bin/env python
"""AQLM Kx8 formats through the fused LUT-GEMV: us per GEMV and GB/s of algorithmic bytes for 1x8 / 2x8 / 4x8 / 8x8 on the
Llama-2-7B shapes, L2-cold (a ring of layers larger than L2), dependent launches in a CUDA graph.  One JSON line per case."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rvv_lut_vq_b200 as r  # noqa: E402
from rvv_lut_vq_b200 import ops  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6553.6)) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6553.6
    for N in (1, 2, 4, 8):
        for I, O in ((4096, 4096), (4096, 11008), (11008, 4096)):
            code_bytes = O * (I // 8) * N
            nbuf = max(4, int(400e6 // code_bytes) + 1)
            g = torch.Generator(device=dev).manual_seed(N)
            layers = [torch.randint(-128, 128, (O, I // 8, N), dtype=torch.int8, device=dev, generator=g) for _ in range(nbuf)]
            cb = torch.randn((N, 256, 1, 8), device=dev, generator=g).half()
            sc = torch.randn((O, 1, 1, 1), device=dev, generator=g).half()
            x = torch.randn((1, I), device=dev, generator=g).half()
            for L in layers[:2]:
                ops.codeKx8_matmat(x, L, cb, sc, None)
            torch.cuda.synchronize()
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for L in layers[:2]:
                    ops.codeKx8_matmat(x, L, cb, sc, None)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=st):
                    for L in layers:
                        ops.codeKx8_matmat(x, L, cb, sc, None)
                for _ in range(3):
                    graph.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                reps = 10
                e0.record()
                for _ in range(reps):
                    graph.replay()
                e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (reps * nbuf)
            alg = code_bytes + N * 256 * 16 + 2 * I + 4 * O
            print(json.dumps({"format": f"{N}x8", "in": I, "out": O, "us_per_gemv": round(us, 2), "algorithmic_bytes": alg, "gbps": round(alg / us / 1e3, 1),
                              "frac_of_hbm_peak": round(alg / us / 1e3 / hbm, 3), "ring": nbuf}), flush=True)
            del layers


if __name__ == "__main__":
    main()
