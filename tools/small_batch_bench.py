#!/usr/bin/env python
"""Batch 1..8 through lutvq_code2x8_matmat_packed on the three Llama-2-7B shapes: us per CALL with the rows of a call grouped into one
launch (workspace ring) and with one launch per row (single region), L2-cold ring of layers, one CUDA graph.  JSON lines."""
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rvv_lut_vq_b200 as r  # noqa: E402
from rvv_lut_vq_b200 import ops, synth  # noqa: E402


def graph_us(st, body, reps, inner):
    with torch.cuda.stream(st):
        body()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            body()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(reps):
            g.replay()
        b.record(st)
        torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / (reps * inner)


def main():
    dev = torch.device("cuda:0")
    lib = r.lib()
    if os.environ.get("GEMV_TUNING"):
        lib.lutvq_debug_set_gemv_tuning(*[int(v) for v in os.environ["GEMV_TUNING"].split(",")])
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    st = torch.cuda.Stream(dev)
    for I, O in ((4096, 11008),) if os.environ.get("GEMV_TUNING") else ((4096, 11008), (4096, 4096), (11008, 4096)):
        NB = 40
        Ls = [synth.make_layer_torch(I, O, dev, seed=k, scales="randn") for k in range(NB)]
        packed = [ops.code2x8_prepack(L["codes"]) for L in Ls]
        need = (int(lib.lutvq_packed_workspace_bytes(O, I)) + 255) // 256 * 256
        ws = torch.zeros(need * 8, dtype=torch.uint8, device=dev)
        for batch in (1, 2, 3, 4, 6, 8):
            x = torch.randn((batch, I), device=dev, dtype=torch.float16)
            y = torch.empty((batch, O), device=dev, dtype=torch.float16)
            out = {"shape": f"{I}x{O}", "batch": batch}
            for name, nbytes in (("grouped_us", need * min(batch, 8)), ("per_row_launch_us", need)):
                def body():
                    s = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
                    for k, L in enumerate(Ls):
                        rc = lib.lutvq_code2x8_matmat_packed(P(x), P(packed[k]), P(L["codebooks"]), P(L["scales"]), None, P(y), batch, O, I, 0, 0,
                                                             P(ws), nbytes, s)
                        assert rc == 0, lib.lutvq_last_error()
                out[name] = round(graph_us(st, body, 5, NB), 2)
            out["vs_batch1"] = None
            out["per_row_us"] = round(out["grouped_us"] / batch, 2)
            print(json.dumps(out), flush=True)
        del Ls, packed
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
