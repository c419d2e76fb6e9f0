"""A few prefill calls (batch from argv, default 16) on 4096x11008 for ncu captures."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rvv_lut_vq_b200 import ops, synth
dev = torch.device('cuda', 0)
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 16
Ls = [synth.make_layer_torch(4096, 11008, dev, seed=k, scales="ones") for k in range(4)]
x = torch.randn((batch, 4096), device=dev, dtype=torch.float16)
for k in range(4):
    ops.code2x8_matmat_dequant(x, Ls[k]["codes"], Ls[k]["codebooks"], Ls[k]["scales"], None)
torch.cuda.synchronize()
