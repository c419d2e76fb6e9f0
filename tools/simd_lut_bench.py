#!/usr/bin/env python
"""lookup-and-add at the reference's own size (bench_avx_lut.cpp main(): 10 000 rows x CB lanes,
100 000 runs, 16-entry int8 table) on one B200, CUDA-event timed, and the scalar oracle on one host core beside it.
Unit = the reference's: lookups per second (GLUT/s)."""
import json, os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
from rvv_lut_vq_b200 import ops
import oracle as orc

n_lut, n_runs = 10000, 100000
for cb in (16, 32, 64):
    rng = np.random.default_rng(cb)
    idx_h = rng.integers(0, 16, size=(n_lut, cb), dtype=np.uint8)
    lut_h = ((np.arange(16) * 10 + 100) & 0xFF).astype(np.uint8).view(np.int8)
    idx, lut = torch.from_numpy(idx_h).cuda(), torch.from_numpy(lut_h).cuda()
    out = torch.zeros(cb, dtype=torch.int16, device="cuda")
    for _ in range(2):
        ops.simd_lut_lookup_add(idx, lut, out, n_runs)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    a.record()
    for _ in range(reps):
        ops.simd_lut_lookup_add(idx, lut, out, n_runs)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    lookups = cb * n_runs * n_lut
    # one host core, a bounded sample of the same walk (200 runs)
    t0 = time.perf_counter()
    orc.c_simd_lut_scalar(idx_h, lut_h, n_runs=200)
    cpu_s = time.perf_counter() - t0
    sm = torch.cuda.get_device_properties(0).multi_processor_count
    print(json.dumps({"op": "simd_lut_lookup_add", "cb_size": cb, "n_lut_per_run": n_lut, "n_runs": n_runs, "ms": round(ms, 3),
                      "glut_per_s": round(lookups / ms / 1e6, 1),
                      "bound": "shared-memory LDS.32, one per lookup: 32 per clock and SM",
                      "frac_of_lds_bound": round(lookups / (ms * 1e-3) / (32 * sm * 1.965e9), 3),
                      "cpu_oracle_scalar_glut_per_s": round(cb * 200 * n_lut / cpu_s / 1e9, 2), "cpu_cores": 1,
                      "reference_published": "1.52 GLUT/s scalar, 38.75 GLUT/s AVX2 (SURVEY 8a a13)"}), flush=True)
