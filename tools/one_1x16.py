#!/usr/bin/env python
"""a few 1x16 GEMV launches on 4096->11008 over distinct layers (for ncu captures)"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rvv_lut_vq_b200 import ops
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(3)
I, O = 4096, 11008
Ls = [torch.randint(-32768, 32768, (O, I // 8, 1), dtype=torch.int16, device=dev, generator=g) for _ in range(12)]
cb = torch.randn((1, 65536, 1, 8), device=dev, generator=g).half()
sc = torch.randn((O, 1, 1, 1), device=dev, generator=g).half()
x = torch.randn((1, I), device=dev, generator=g).half()
for c in Ls:
    y = ops.code1x16_matmat(x, c, cb, sc, None)
torch.cuda.synchronize()
print("ok", float(y.float().abs().max()))
