#!/usr/bin/env python
"""The reference's own CUDA kernels for the OTHER formats (1x16: Code1x16MatVec, the default of benchmark_rq_gemm_gpu.py; 1x8:
CodeKx8MatVec<K=1>) next to ours on the same B200: us per GEMV on the Llama-2-7B shapes, ring of L2-cold layers, one CUDA
graph, dependent launches; outputs compared.  JSON lines.  (2x8: tools/ref_cuda_bench.py / bench.py kernel_to_beat.)"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from ref_cuda_bench import load_reference, time_graph  # noqa: E402
from rvv_lut_vq_b200 import ops  # noqa: E402


def main():
    ref, why = load_reference()
    if ref is None:
        print(json.dumps({"unavailable": why}))
        return
    dev = torch.device("cuda:0")
    st = torch.cuda.Stream(dev)
    g = torch.Generator(device=dev).manual_seed(3)
    with torch.cuda.stream(st):
        for fmt in ("1x16", "1x8"):
            for I, O in ((4096, 4096), (4096, 11008), (11008, 4096)):
                if fmt == "1x16":
                    code_bytes, mk = O * (I // 8) * 2, lambda: torch.randint(-32768, 32768, (O, I // 8, 1), dtype=torch.int16, device=dev, generator=g)
                    cb = torch.randn((1, 65536, 1, 8), device=dev, generator=g).half()
                    ours, theirs = ops.code1x16_matmat, ref.code1x16_matmat
                else:
                    code_bytes, mk = O * (I // 8), lambda: torch.randint(-128, 128, (O, I // 8, 1), dtype=torch.int8, device=dev, generator=g)
                    cb = torch.randn((1, 256, 1, 8), device=dev, generator=g).half()
                    ours, theirs = ops.code1x8_matmat, ref.code1x8_matmat
                nb = min(64, max(8, int(300e6 // code_bytes) + 1))
                Ls = [mk() for _ in range(nb)]
                sc = torch.randn((O, 1, 1, 1), device=dev, generator=g).half()
                x = torch.randn((1, I), device=dev, generator=g).half()
                res = {}

                def body(fn):
                    def b():
                        for k, c in enumerate(Ls):
                            res[k] = fn(x, c, cb, sc, None)
                    return b
                us_ref = time_graph(torch, st, body(theirs), 5) / nb
                y_ref = res[0].clone()
                us_our = time_graph(torch, st, body(ours), 5) / nb
                y_our = res[0].clone()
                extra = {}
                if fmt == "1x16" and os.environ.get("LUTVQ_1X16_MODES"):
                    import rvv_lut_vq_b200 as r
                    for mode, name in ((1, "ours_hfma2_us"), (2, "ours_gather_only_us")):
                        r.lib().lutvq_debug_set_1x16_mode(mode)
                        extra[name] = round(time_graph(torch, st, body(ours), 5) / nb, 2)
                    r.lib().lutvq_debug_set_1x16_mode(0)
                den = float(y_ref.float().abs().max())
                print(json.dumps({"format": fmt, "shape": f"{I}->{O}", "ring": nb, "reference_us": round(us_ref, 2), "ours_us": round(us_our, 2),
                                  "speedup": round(us_ref / us_our, 2), "max_abs_diff_over_max": float((y_our.float() - y_ref.float()).abs().max()) / den, **extra}), flush=True)
                # dequantize (+ scales) and a batch-4 call of the same format
                if (I, O) == (4096, 11008):
                    dq_ours = ops.code1x16_dequant if fmt == "1x16" else ops.code1x8_dequant
                    dq_ref = ref.code1x16_dequant if fmt == "1x16" else ref.code1x8_dequant

                    def dq(fn):
                        def b():
                            for k, c in enumerate(Ls[:8]):
                                res[k] = fn(c, cb, sc)
                        return b
                    t_ref = time_graph(torch, st, dq(dq_ref), 5) / 8
                    w_ref = res[0].clone()
                    t_our = time_graph(torch, st, dq(dq_ours), 5) / 8
                    print(json.dumps({"format": fmt, "op": "dequant", "shape": f"{I}x{O}", "reference_us": round(t_ref, 2), "ours_us": round(t_our, 2),
                                      "speedup": round(t_ref / t_our, 2), "bit_identical": bool(torch.equal(res[0].view(torch.int16), w_ref.view(torch.int16)))}), flush=True)
                    x4 = torch.randn((4, I), device=dev, generator=g).half()

                    def b4(fn):
                        def b():
                            for k, c in enumerate(Ls):
                                res[k] = fn(x4, c, cb, sc, None)
                        return b
                    t_ref = time_graph(torch, st, b4(theirs), 5) / nb
                    y4 = res[0].clone()
                    t_our = time_graph(torch, st, b4(ours), 5) / nb
                    print(json.dumps({"format": fmt, "op": "matmat batch 4", "shape": f"{I}->{O}", "reference_us": round(t_ref, 2), "ours_us": round(t_our, 2),
                                      "speedup": round(t_ref / t_our, 2),
                                      "max_abs_diff_over_max": float((res[0].float() - y4.float()).abs().max()) / float(y4.float().abs().max())}), flush=True)
                del Ls, res


if __name__ == "__main__":
    main()
