#!/usr/bin/env python
"""T-MAC bit-serial GEMV (lutvq_tmac_preprocessor_int8 + lutvq_tmac_qgemm_lut_int8) on the GPU next to the CPU port of the same
call structure (oracle/tmac_oracle.c, one host thread, bit-identical results): us per GEMV at LLM sizes.  JSON lines."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
from rvv_lut_vq_b200 import ops  # noqa: E402
import oracle as orc  # noqa: E402


def main():
    cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    for M, K, bits, bm in ((4096, 4096, 4, 512), (4096, 4096, 2, 512), (11008, 4096, 4, 512), (4096, 11008, 4, 512)):
        rng = np.random.default_rng(M + K + bits)
        B = rng.standard_normal(K).astype(np.float32)
        tiles, kt_n = M * bits // bm, K // 32
        A = rng.integers(0, 256, size=(tiles, kt_n, bm // 2 * 8), dtype=np.uint8)
        Sc = (rng.standard_normal((tiles, kt_n, bm // bits)) * 0.1).astype(np.float32)
        dB, dA, dS = cu(B), cu(A), cu(Sc)
        for _ in range(3):
            q, ls, lb = ops.tmac_preprocessor_int8(dB)
            C = ops.tmac_qgemm_lut_int8(M, K, bits, bm, dA, q, dS, ls, lb)
        torch.cuda.synchronize()
        # one CUDA graph per op, 20 calls each: the eager wrappers (three torch.empty per call) cost more host time than the kernels run
        st = torch.cuda.Stream()
        reps = 20
        times = []
        with torch.cuda.stream(st):
            for body in (lambda: ops.tmac_preprocessor_int8(dB), lambda: ops.tmac_qgemm_lut_int8(M, K, bits, bm, dA, q, dS, ls, lb)):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=st):
                    for _ in range(reps):
                        out = body()
                g.replay()
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(st)
                for _ in range(5):
                    g.replay()
                b.record(st)
                torch.cuda.synchronize()
                times.append(a.elapsed_time(b) * 1e3 / (5 * reps))
        C = ops.tmac_qgemm_lut_int8(M, K, bits, bm, dA, q, dS, ls, lb)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        q0, ls0, lb0 = orc.c_tmac_preprocessor_int8(B)
        want = orc.c_tmac_qgemm_lut_int8(M, K, bits, bm, A, q0, Sc, ls0, lb0)
        cpu_us = (time.perf_counter() - t0) * 1e6
        same = bool(np.array_equal(C.cpu().numpy().view(np.uint32), want.view(np.uint32)))
        nbytes = A.size + Sc.size * 4 + K * 4 + M * 4
        us_q = times[1]
        print(json.dumps({"M": M, "K": K, "bits": bits, "bm": bm, "preprocessor_us": round(times[0], 1), "qgemm_us": round(us_q, 1),
                          "algorithmic_bytes": nbytes, "gbps": round(nbytes / us_q / 1e3, 1), "cpu_port_1thread_us": round(cpu_us, 1), "bit_identical": same}), flush=True)


if __name__ == "__main__":
    main()
