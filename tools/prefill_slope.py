"""Prefill kernel time against the k extent (fixed cost vs per-k-block cost), per debug mode."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rvv_lut_vq_b200 import ops, synth, _lib
from prefill_bench import timed
dev = torch.device('cuda', 0)
for I in (2048, 8192):
    nl = max(4, min(24, (300 << 20) // (11008 * I // 4)))
    Ls = [synth.make_layer_torch(I, 11008, dev, seed=k, scales="ones") for k in range(nl)]
    x = torch.randn((16, I), device=dev, dtype=torch.float16)
    out = []
    for mode in (55, 119, 71):
        _lib.lib().lutvq_debug_set_gemm_mode(mode)
        k = [0]
        def f():
            L = Ls[k[0] % nl]; k[0] += 1
            ops.code2x8_matmat_dequant(x, L["codes"], L["codebooks"], L["scales"], None)
        out.append(round(timed(f, 48), 1))
    print("I", I, "kblocks", I // 64, "us mode0/2/7", out, flush=True)
_lib.lib().lutvq_debug_set_gemm_mode(0)
