"""Prefill kernel with parts switched off (lutvq_debug_set_gemm_mode: 1 = no MMAs, 2 = no dequantization, 4 = no
activation copies, 8 = 128-row tiles, 16 = no proxy fence, 128 = 1 k block per stage, 256 = 2, 2048 = no programmatic dependent launch); timing only, results are wrong in every mode but 0.  Graph-captured, ring of 24 layers."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rvv_lut_vq_b200 import ops, synth, _lib
from prefill_bench import timed
dev = torch.device('cuda', 0)
Ls = [synth.make_layer_torch(4096, 11008, dev, seed=k, scales="ones") for k in range(24)]
MODES = [int(a) for a in sys.argv[1:]] or [0]
for batch in (16, 64, 128, 256):
    x = torch.randn((batch, 4096), device=dev, dtype=torch.float16)
    for mode in MODES:
        _lib.lib().lutvq_debug_set_gemm_mode(mode)
        k = [0]
        def f():
            L = Ls[k[0] % 24]; k[0] += 1
            ops.code2x8_matmat_dequant_fused(x, L["codes"], L["codebooks"], L["scales"], None)
        print(batch, "mode", mode, round(timed(f, 48), 1), "us", flush=True)
_lib.lib().lutvq_debug_set_gemm_mode(0)
