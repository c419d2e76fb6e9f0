"""Traffic counters: the reference's dataflow simulator restated in closed-loop form, and the same
accounting for the memory levels of a B200 running this library's fused LUT-GEMV.

Reference: ``vq_dataflow_sim/vq.py:68-267`` (``VQ.dataflow_sim``) walks seven loop orders over (O = output rows, D = batch
rows, M = subvectors, N = codebooks) with tile sizes derived from a vector length, and counts the BITS that cross the
off-chip boundary per operand class: codewords (``BW`` = log2 K bits each), LUT entries, partial sums, inputs, codebooks
(16 bits each).  ``reference_dataflow_traffic`` replays exactly those loop nests on sizes only (no arrays), so it is cheap at
any shape; tests/test_host_logic.py pins it against rows produced by the reference's own code (tests/golden/dataflow_sim_ref.csv).

``gemv_traffic_model`` is the counterpart for OUR dataflow on the GPU -- the reference's "VeLU" order (LUT built on chip and
never spilled, codewords streamed exactly once) with the k dimension split across CTAs -- per memory level:
HBM (= the algorithmic bytes of DESIGN.md), L2 -> SM, L2 atomics, shared memory.  bench.py / DESIGN.md compare its HBM figure
with the DRAM bytes ncu measured for the kernel (profiles/traffic.json)."""
from __future__ import annotations

import math
from typing import Dict, Tuple


def _tiling(cfg: Tuple[int, int, int, int], VLEN: int):
    """vq.py:69-82 get_tiling_parms"""
    D, M, N, K = cfg
    d = D // M
    BW = math.log2(K)
    n_codebook_tile = N
    while VLEN // 16 // n_codebook_tile // K == 0:
        n_codebook_tile //= 2
    n_cluster_tile = VLEN // 16 // d // n_codebook_tile
    LMUL = 16 // BW
    n_subvec_tile = int(LMUL * VLEN) // 16 // n_codebook_tile // K
    n_cw_tile = int(K * 16 // BW)
    return dict(BW=BW, d=d, n_codebook_tile=n_codebook_tile, n_cluster_tile=n_cluster_tile, LMUL=LMUL, n_subvec_tile=n_subvec_tile,
                n_cw_tile=n_cw_tile)


def reference_dataflow_traffic(cfg: Tuple[int, int, int, int], d_in: int, d_out: int, dataflow: str, VLEN: int = 4096) -> Dict[str, float]:
    """Off-chip traffic in MB (bits / 8 / 2**20, as sim_dataflow.py:96-100 reports it) of VQ.dataflow_sim for cfg = (D, M, N, K).
    Keys: codeword, codebook, lut, inp, psum, total, plus the tile sizes."""
    D, M, N, K = cfg
    t = _tiling(cfg, VLEN)
    BW, d = t["BW"], t["d"]
    mt, nt, jt = t["n_subvec_tile"], t["n_codebook_tile"], t["n_cw_tile"]
    cnt = dict(codeword=0.0, codebook=0.0, lut=0.0, inp=0.0, psum=0.0)
    rng = lambda n, step: [(s, min(step, n - s)) for s in range(0, n, step)] if step > 0 else []
    ms, ns, js = rng(M, mt), rng(N, nt), rng(d_out, jt)

    def compute_lut():                                   # vq.py:98-102 _compute_lut (one input row)
        cnt["inp"] += D * 16
        cnt["codebook"] += M * N * K * d * 16

    def load_lut_full():                                 # _load_lut(self.lut): the whole (M, N, K) table of one row
        cnt["lut"] += M * N * K * 16

    def load_lut(m, n):
        cnt["lut"] += m * n * K * 16

    def load_cw(j, m, n):
        cnt["codeword"] += j * m * n * BW

    def load_psum(j):
        cnt["psum"] += j * 16

    spill_lut = dataflow in ("OMND", "DMNO", "MNOD", "ODMN", "DOMN")
    if dataflow in ("OMND", "DMNO", "MNOD", "MNDO", "ODMN", "DOMN"):
        for _ in range(d_in):
            compute_lut()
            if spill_lut:
                load_lut_full()                          # the LUT of every input row is written out once (MNDO keeps it: vq.py:151-152)
    if dataflow == "OMND":
        for _, j in js:
            for _, m in ms:
                for _, n in ns:
                    load_cw(j, m, n)
                    for _ in range(d_in):
                        load_psum(j)
                        load_lut(m, n)
    elif dataflow == "DMNO":
        for _ in range(d_in):
            for _, m in ms:
                for _, n in ns:
                    load_lut(m, n)
                    for _, j in js:
                        load_psum(j)
                        load_cw(j, m, n)
    elif dataflow == "MNOD":
        for _, m in ms:
            for _, n in ns:
                for _, j in js:
                    load_cw(j, m, n)
                    for _ in range(d_in):
                        load_psum(j)
                        load_lut(m, n)
    elif dataflow == "MNDO":
        for _, m in ms:
            for _, n in ns:
                for _ in range(d_in):
                    load_lut(m, n)
                    for _, j in js:
                        load_psum(j)
                        load_cw(j, m, n)
    elif dataflow == "ODMN":
        for _, j in js:
            for _ in range(d_in):
                load_psum(j)
                for _, m in ms:
                    for _, n in ns:
                        load_lut(m, n)
                        load_cw(j, m, n)
    elif dataflow == "DOMN":
        for _ in range(d_in):
            for _, j in js:
                load_psum(j)
                for _, m in ms:
                    for _, n in ns:
                        load_lut(m, n)
                        load_cw(j, m, n)
    elif dataflow == "VeLU":
        for _ in range(d_in):
            compute_lut()                                # the LUT stays on chip: never counted as off-chip traffic
            for _, m in ms:
                for _, n in ns:
                    for _, j in js:
                        load_psum(j)
                        load_cw(j, m, n)
    else:
        raise NotImplementedError(f"Dataflow {dataflow} not implemented")      # the '-compact' orders are empty in the reference too
    out = {k: v / 1024 ** 2 / 8 for k, v in cnt.items()}
    out["total"] = sum(out.values())
    out.update(N_tile=nt, M_tile=mt, CW_tile=jt)
    return out


def gemv_traffic_model(in_features: int, out_features: int, n_sm: int = 148, workers_per_sm: int = 1, slice_tables: int = 64,
                       lut_entry_bytes: int = 4) -> Dict[str, float]:
    """Bytes per memory level of ONE fused LUT-GEMV launch (AQLM 2x8, batch 1) of this library.
    hbm       every operand once: codes + codebooks + x + scales + y  (= DESIGN.md's algorithmic bytes)
    l2_to_sm  what the SMs pull through L2: the code stream once, the 8 KiB codebook once per CTA, each CTA's slice of x,
              the scales of the finishing rows
    l2_atomic split-k traffic: one 64-bit atomic (request + returned word) per (row, k-slice)
    smem_st   LUT builds: one table set per CTA and k-slice it touches (64 KiB each)
    smem_ld   one entry per code byte
    The LUT itself (2 x 256 x G entries) never leaves the chip -- the reference's 'VeLU' order."""
    G = in_features // 8
    n_k = -(-(2 * G) // slice_tables)
    ctas = n_sm * workers_per_sm
    codes = out_features * G * 2
    codebook = 2 * 256 * 8 * 2
    hbm = codes + codebook + 2 * in_features + 2 * out_features + 2 * out_features
    builds = max(ctas, n_k)                                          # aligned plan: every CTA builds exactly one slice
    slice_x = slice_tables // 2 * 16
    return {
        "hbm": float(hbm),
        "l2_to_sm": float(codes + ctas * codebook + builds * slice_x + 2 * out_features),
        "l2_atomic": float(out_features * n_k * 16),
        "smem_st": float(builds * slice_tables * 256 * lut_entry_bytes),
        "smem_ld": float(codes * lut_entry_bytes),
        "n_k": n_k, "ctas": ctas,
    }
