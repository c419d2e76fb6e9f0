#!/usr/bin/env python
"""(Needs the instrumented build: make -C rvv-lut-vq_b200/csrc clean all EXTRA=-DLUTVQ_GEMM_TIMELINE.)
Where a CTA of the prefill kernel spends its life: %globaltimer stamps of thread 0 (lutvq_debug_set_gemm_stamps) for a run of
back-to-back launches on 4096x11008; prints, per batch size, the average over CTAs and launches of each interval in us.
stamps: 0 entry, 1 set-up done, 2 dependency wait passed, 3 first code chunk's stages handed over, 4 producer loop done,
5 accumulator complete, 6 epilogue done; slots 8..12: cycle totals of producer thread 0 and of the MMA thread."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rvv_lut_vq_b200 import ops, synth, _lib

dev = torch.device("cuda", 0)
NB, LAUNCHES = 24, 48
I_F, O_F = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4096, 11008)
Ls = [synth.make_layer_torch(I_F, O_F, dev, seed=k, scales="ones") for k in range(NB)]
lib = _lib.lib()
NCTA = int(sys.argv[3]) if len(sys.argv) > 3 else 138                 # CTAs per launch (138 for 4096x11008, 128 for the split-k shapes)
NAMES = ["set-up", "dependency wait", "first chunk", "rest of the loop", "wait for the accumulator", "epilogue"]
for batch in (16, 64, 256):
    x = torch.randn((batch, I_F), device=dev, dtype=torch.float16)
    bufs = [torch.zeros((160, 16), dtype=torch.int64, device=dev) for _ in range(LAUNCHES)]
    for _ in range(2):
        for k in range(LAUNCHES):
            lib.lutvq_debug_set_gemm_stamps(ctypes.c_void_p(bufs[k].data_ptr()))
            L = Ls[k % NB]
            ops.code2x8_matmat_dequant_fused(x, L["codes"], L["codebooks"], L["scales"], None)
    lib.lutvq_debug_set_gemm_stamps(None)
    torch.cuda.synchronize()
    assert int(bufs[-1][0, 0]) != 0, "no stamps: liblutvq.so was built without -DLUTVQ_GEMM_TIMELINE"
    t = torch.stack(bufs).double().cpu()[8:, :NCTA, :7]                     # [launch][CTA][stamp], first launches dropped
    iv = (t[:, :, 1:] - t[:, :, :-1]).mean(dim=(0, 1)) / 1e3
    life = (t[:, :, 6] - t[:, :, 0]).mean() / 1e3
    start = t[:, :, 0].min(dim=1).values
    period = ((start[1:] - start[:-1]).mean() / 1e3).item()
    spread = ((t[:, :, 0].max(dim=1).values - start).mean() / 1e3).item()
    end_spread = ((t[:, :, 6].max(dim=1).values - t[:, :, 6].min(dim=1).values).mean() / 1e3).item()
    print(f"batch {batch}: launch period {period:.2f} us (eager launches), CTA life {life:.2f} us, entry spread {spread:.2f}, exit spread {end_spread:.2f}")
    print("   " + ", ".join(f"{n} {v:.2f}" for n, v in zip(NAMES, iv.tolist())), flush=True)
    tot = torch.stack(bufs).double().cpu()[8:, :NCTA, 8:13].mean(dim=(0, 1)) / 1965.0      # cycles -> us at 1965 MHz
    print("   thread 0 of the producers: waiting for an empty stage %.2f us, code staging %.2f, waiting for its tensor-memory stores %.2f; "
          "MMA thread: waiting for a full stage %.2f of its %.2f us loop" % tuple(tot.tolist()), flush=True)
