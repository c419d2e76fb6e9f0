#!/usr/bin/env python
"""The prefill op's two routes on the other Llama-2-7B shapes (split-k plans): us per call, graph-captured ring of L2-cold layers."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rvv_lut_vq_b200 import ops, synth
from prefill_bench import timed
dev = torch.device("cuda", 0)
for (I, O) in [(4096, 4096), (11008, 4096), (4096, 12288), (4096, 22016)]:
    NB = max(4, min(24, int(300e6 / (I * O / 4))))
    Ls = [synth.make_layer_torch(I, O, dev, seed=k, scales="ones") for k in range(NB)]
    for batch in (16, 64, 256):
        x = torch.randn((batch, I), device=dev, dtype=torch.float16)
        row = {"shape": f"{I}x{O}", "batch": batch, "layers_in_ring": NB}
        for name, fn in (("fused_us", ops.code2x8_matmat_dequant_fused), ("two_step_us", ops.code2x8_matmat_dequant_two_step)):
            k = [0]
            def f():
                L = Ls[k[0] % NB]; k[0] += 1
                fn(x, L["codes"], L["codebooks"], L["scales"], None)
            row[name] = round(timed(f, 2 * NB), 2)
        print(json.dumps(row), flush=True)
    del Ls
    torch.cuda.empty_cache()
