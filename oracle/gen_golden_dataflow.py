#!/usr/bin/env python
"""Golden rows for the dataflow / traffic counter -- TEST INFRASTRUCTURE.

Runs the reference's own VQ.dataflow_sim (vq_dataflow_sim/vq.py:68-267) here for the 7 implemented loop orders x the 4
configurations of vq_dataflow_sim/sim_dataflow.py:33-52 and freezes the counters (MB, as sim_dataflow.py:96-100 converts
them) under tests/golden/dataflow_sim_ref.csv.  The table the reference ships (vq_dataflow_sim/dataflow_sim.csv) is
compared row by row; it differs from the code as it stands only in the VeLU psum column of RQ / AQLM_GEMV (an older loop nest).
Only runnable in the build container (a few minutes: the constructors dequantize in Python loops); the fixture travels."""
import contextlib
import csv
import io
import os
import sys

import numpy as np

REF = "/root/reference/vq_dataflow_sim"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden", "dataflow_sim_ref.csv")
CONFIGS = {"TMAC_GEMV": ((4096, 1024, 4, 16), 4096, "TMAC"), "PQ": ((128, 8, 1, 256), 1024, None),
           "RQ": ((128, 1, 4, 256), 1024, None), "AQLM_GEMV": ((4096, 256, 2, 256), 4096, None)}
FLOWS = ["OMND", "ODMN", "DMNO", "MNOD", "MNDO", "DOMN", "VeLU"]
KEYS = ["codeword", "codebook", "lut", "inp", "psum"]


def main():
    sys.path.insert(0, REF)
    with contextlib.redirect_stdout(io.StringIO()):
        import vq
    shipped = {(r["Dataflow"], r["VQ_type"]): r for r in csv.DictReader(open(os.path.join(REF, "dataflow_sim.csv")))}
    rows = []
    for name, (cfg, d_out, vq_type) in CONFIGS.items():
        with contextlib.redirect_stdout(io.StringIO()):
            m = vq.VQ(cfg, d_out=d_out, vq_type=vq_type)
        x = np.random.randn(1, cfg[0]).astype(np.float16)
        for flow in FLOWS:
            for k in m.perf_cnt:
                m.perf_cnt[k] = 0
            with contextlib.redirect_stdout(io.StringIO()):
                m.dataflow_sim(x, flow)
            mb = {k: float(m.perf_cnt["offchip_access_" + k]) / 1024 ** 2 / 8 for k in KEYS}
            rows.append(dict(Dataflow=flow, VQ_type=name, Din=1, Dout=d_out, VLEN=m.VLEN, D=cfg[0], M=cfg[1], N=cfg[2], K=cfg[3],
                             N_tile=m.n_codebook_tile, M_tile=m.n_subvec_tile, CW_tile=m.n_cw_tile, **{k + "_MB": v for k, v in mb.items()}))
            s = shipped.get((flow, name))
            diff = [k for k in KEYS if s and abs(float(s[f"Offchip_access_{k}_MB"]) - mb[k]) > 1e-12]
            print(name, flow, mb, "DIFFERS from the shipped csv in " + ",".join(diff) if diff else "== shipped csv")
    with open(OUT, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
        w.writeheader()
        w.writerows(rows)
    print("wrote", OUT, len(rows), "rows")


if __name__ == "__main__":
    main()
