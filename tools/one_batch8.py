#!/usr/bin/env python
"""a few batch-8 grouped GEMV launches on 4096->11008 over distinct layers (for ncu captures of the gather loop)"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import rvv_lut_vq_b200 as r
from rvv_lut_vq_b200 import ops, synth
dev = torch.device("cuda:0")
I, O, B = 4096, 11008, 8
Ls = [synth.make_layer_torch(I, O, dev, seed=k, scales="randn") for k in range(12)]
pk = [ops.code2x8_prepack(L["codes"]) for L in Ls]
x = torch.randn((B, I), device=dev, dtype=torch.float16)
for k, L in enumerate(Ls):
    y = ops.code2x8_matmat_packed(x, pk[k], O, L["codebooks"], L["scales"], None)
torch.cuda.synchronize()
print("ok", float(y.float().abs().max()))
