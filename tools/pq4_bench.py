#!/usr/bin/env python
"""PQ4 fast-scan accumulation: GB/s of packed codes on the GPU (L2-cold: the code image is larger than L2) next to the
reference's own AVX2 kernel on one host core (oracle/_ref/libpq4_ref.so, when it travelled).  One JSON line per case."""
import ctypes
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import rvv_lut_vq_b200 as r  # noqa: E402
from rvv_lut_vq_b200 import pq4  # noqa: E402


def main():
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = float(peaks.get("hbm_gbs", 6553.6))
    ref_path = os.path.join(ROOT, "oracle", "_ref", "libpq4_ref.so")
    ref = ctypes.CDLL(ref_path) if os.path.exists(ref_path) else None
    dev = torch.device("cuda:0")
    cases = [(1, 32, 32, 1 << 24), (2, 32, 32, 1 << 24), (4, 32, 32, 1 << 24), (1, 256, 32, 1 << 24), (1, 32, 64, 1 << 23), (1, 32, 16, 1 << 25),
                             (4, 32, 64, 1 << 23)]
    if os.environ.get("PQ4_BENCH_CASES"):
        cases = [cases[int(k)] for k in os.environ["PQ4_BENCH_CASES"].split(",")]
    for nq, bbs, nsq, nb in cases:
        g = torch.Generator(device=dev).manual_seed(1)
        blocks = torch.randint(0, 256, (nb * nsq // 2,), dtype=torch.uint8, device=dev, generator=g)
        lut = torch.randint(0, 256, (nq * nsq * 16,), dtype=torch.uint8, device=dev, generator=g)
        accu = torch.empty((nq, nb), dtype=torch.uint16, device=dev)
        for _ in range(3):
            pq4.pq4_accumulate_loop(nq, nb, bbs, nsq, blocks, lut, accu)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            pq4.pq4_accumulate_loop(nq, nb, bbs, nsq, blocks, lut, accu)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps
        bytes_alg = nb * nsq // 2 + nq * nb * 2 + nq * nsq * 16
        out = {"nq": nq, "bbs": bbs, "nsq": nsq, "nb": nb, "us": round(us, 1), "algorithmic_bytes": bytes_alg, "gbps": round(bytes_alg / us / 1e3, 1),
               "frac_of_hbm_peak": round(bytes_alg / us / 1e3 / hbm, 3), "lookups_per_ns": round(nq * nb * nsq / us / 1e3, 1)}
        if ref is not None and nq * 1000 + bbs // 32 in (1001, 2001, 4001, 1002, 1003, 1004, 1005, 2002, 3001):
            n_cpu = 1 << 20                                              # a bounded sample of the same database
            hb = blocks[: n_cpu * nsq // 2].cpu().numpy()
            hl = lut.cpu().numpy()
            raw = np.zeros(hb.size + 64, np.uint8); o = (-raw.ctypes.data) % 32; ab = raw[o:o + hb.size]; ab[:] = hb
            rawl = np.zeros(hl.size + 64, np.uint8); o = (-rawl.ctypes.data) % 32; al = rawl[o:o + hl.size]; al[:] = hl
            acc = np.zeros((nq, n_cpu), np.uint16)
            P, Z, I = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int
            ref.ref_pq4_accumulate_loop.argtypes = [I, Z, I, I, P, P, P, I]
            ref.ref_pq4_accumulate_loop(nq, n_cpu, bbs, nsq, ab.ctypes.data_as(P), al.ctypes.data_as(P), acc.ctypes.data_as(P), -1)
            t0 = time.perf_counter()
            for _ in range(3):
                ref.ref_pq4_accumulate_loop(nq, n_cpu, bbs, nsq, ab.ctypes.data_as(P), al.ctypes.data_as(P), acc.ctypes.data_as(P), -1)
            cpu_us = (time.perf_counter() - t0) / 3 * 1e6
            same = bool(np.array_equal(acc, accu[:, :n_cpu].cpu().numpy()))
            out.update(reference_avx2_1core_gbps=round((n_cpu * nsq // 2 + nq * n_cpu * 2) / cpu_us / 1e3, 2), reference_sample_vectors=n_cpu, bit_identical_to_reference=same)
        r._lib.lib().lutvq_debug_set_pq4_mode(1)                          # the PRMT kernel on the same inputs
        for _ in range(2):
            pq4.pq4_accumulate_loop(nq, nb, bbs, nsq, blocks, lut, accu)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            pq4.pq4_accumulate_loop(nq, nb, bbs, nsq, blocks, lut, accu)
        e1.record()
        torch.cuda.synchronize()
        r._lib.lib().lutvq_debug_set_pq4_mode(0)
        out["prmt_kernel_us"] = round(e0.elapsed_time(e1) * 1e3 / reps, 1)
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
