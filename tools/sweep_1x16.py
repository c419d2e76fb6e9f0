#!/usr/bin/env python
"""1x16 GEMV launch-plan sweep (experiment): us per GEMV for split x CTAs-per-SM on the three 7B shapes, L2-cold ring, CUDA graph."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
from ref_cuda_bench import time_graph  # noqa: E402
import rvv_lut_vq_b200 as r  # noqa: E402
from rvv_lut_vq_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
st = torch.cuda.Stream(dev)
g = torch.Generator(device=dev).manual_seed(3)
lib = r.lib()
with torch.cuda.stream(st):
    for I, O in ((4096, 4096), (4096, 11008), (11008, 4096)):
        nb = min(64, max(8, int(300e6 // (O * (I // 8) * 2)) + 1))
        Ls = [torch.randint(-32768, 32768, (O, I // 8, 1), dtype=torch.int16, device=dev, generator=g) for _ in range(nb)]
        cb = torch.randn((1, 65536, 1, 8), device=dev, generator=g).half()
        sc = torch.randn((O, 1, 1, 1), device=dev, generator=g).half()
        x = torch.randn((1, I), device=dev, generator=g).half()
        res = {}
        def body():
            for k, c in enumerate(Ls):
                res[k] = ops.code1x16_matmat(x, c, cb, sc, None)
        row = {"shape": f"{I}->{O}"}
        for rep in range(2):
            for split in (1, 2, 4):
                for per_sm in (1, 2, 3, 4, 6, 8):
                    lib.lutvq_debug_set_1x16_plan(split, per_sm)
                    row[f"s{split}p{per_sm}r{rep}"] = round(time_graph(torch, st, body, 4) / nb, 2)
        lib.lutvq_debug_set_1x16_plan(0, 0)
        row["auto"] = round(time_graph(torch, st, body, 4) / nb, 2)
        print(json.dumps(row), flush=True)
        del Ls, res
