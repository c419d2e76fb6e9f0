#!/usr/bin/env python
"""bench.py -- LUT-VQ GEMV throughput on B200 (BASELINE.json metric: "LUT-VQ GEMV us & HBM GB/s (AQLM 2x8,
4096x11008)"), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (N=1, BASELINE configs[1]): Llama-2-7B gate_proj shape 4096->11008, AQLM 2x8 codes, batch 1.
One STEP = one LUT-GEMV over each of NBUF distinct layers (random codes/codebooks/scales; NBUF = 64 on one GPU = the
gate/up projections of one Llama-2-7B token, 24 per rank when sharded), whose code matrices total > 2x the 126 MB L2,
so every launch streams its codes from HBM (L2-cold by construction).
  value   whole-job GB/s on the algorithmic byte stream (O*(I/8)*2 + 8192 + 2I + 4O per GEMV),
          inputs resident in HBM, the NBUF launches replayed as one CUDA graph, timed with CUDA events;
  e2e     the same metric through the reference-facing op (ops.code2x8_matmat) with HOST buffers: per GEMV the
          activation row is copied from pinned host memory and the output row copied back, inside the timed region;
  roofline / cpu_baseline / clocks / sweep: see DESIGN.md "Measurement".
N>1 (torchrun, one rank per GPU): the layer is sharded by OUTPUT CHANNEL (each rank owns an 11008-row shard of
an N*11008-row layer), every GEMV's output slice is all-gathered into every rank's gather slot by the push exchange
(one exchange launch per step of 24 independent GEMVs by default, LUTVQ_BENCH_GROUP; weak scaling).
--impl reference: the reference's CPU LUT-GEMV (oracle port of numba_gemv_lut_, all host threads) on the same config.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

IN_F, OUT_F = 4096, 11008           # headline shape (gate_proj), BASELINE.json metric
# GEMVs per step.  One GPU: 64 = the gate_proj + up_proj launches of one Llama-2-7B decode token (32 layers x 2, all of the
# headline shape), 64 x 11.27 MB = 721 MB of codes per step.  Sharded (N > 1): 24 (the gather-slot ring holds 24 slots),
# 270 MB per rank and step.  Both are > 2x L2.  The step is one CUDA graph; a graph boundary costs one programmatic-
# dependent-launch overlap (~7 us): 8 % of a 24-launch step, 3 % of a 64-launch step (measured 2835 / 2989 / 3033 GB/s
# at 24 / 72 / 144 launches per step), so `value` is the steady-state stream plus that stated share of boundary cost.
NBUF = int(os.environ.get("LUTVQ_BENCH_NBUF", "0")) or (64 if int(os.environ.get("WORLD_SIZE", "1")) == 1 else 24)
METRIC = "lutvq_gemv_hbm_gbps"
UNIT = "GB/s"


def algorithmic_bytes(i: int, o: int) -> int:
    return o * (i // 8) * 2 + 2 * 256 * 8 * 2 + 2 * i + 4 * o


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons while the timed region runs"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """The reference's CPU implementation of the path (numba_gemv_lut_ semantics, aqlm numba_kernel.py:38-48),
    timed as the oracle port on all host threads; each step = CPU_BUFS GEMVs of the headline shape."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as orc
    CPU_BUFS = 4
    rng = np.random.default_rng(0)
    G = IN_F // 8
    layers = [(rng.integers(0, 256, size=(OUT_F, G, 2), dtype=np.uint8).view(np.int8),
               rng.standard_normal((2, 256, 8)).astype(np.float16).view(np.uint16),
               np.ones(OUT_F, dtype=np.float16).view(np.uint16)) for _ in range(CPU_BUFS)]
    x = rng.standard_normal(IN_F).astype(np.float16).view(np.uint16)
    f = orc.RowMajorGemv(OUT_F, G)
    orc.calibrate_threads(lambda: f(x, *layers[0]))
    step = lambda: [f(x, c, cb, sc) for (c, cb, sc) in layers]
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    ms = dt / args.steps * 1e3
    val = CPU_BUFS * algorithmic_bytes(IN_F, OUT_F) / (ms * 1e-3) / 1e9
    cores = orc.num_threads()
    sample = f"{args.steps} steps x {CPU_BUFS} GEMVs of {IN_F}x{OUT_F} (distinct code matrices), {cores} OpenMP threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"AQLM 2x8 LUT-GEMV {IN_F}x{OUT_F} batch 1 (Llama-2-7B gate_proj), CPU LUT path", "gemvs_per_step": CPU_BUFS},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "us_per_gemv": ms * 1e3 / CPU_BUFS,
    }))


def cpu_baseline(seconds: float = 12.0):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as orc
    rng = np.random.default_rng(0)
    G = IN_F // 8
    bufs = 4
    layers = [(rng.integers(0, 256, size=(OUT_F, G, 2), dtype=np.uint8).view(np.int8),
               rng.standard_normal((2, 256, 8)).astype(np.float16).view(np.uint16)) for _ in range(bufs)]
    x = rng.standard_normal(IN_F).astype(np.float16).view(np.uint16)
    f = orc.RowMajorGemv(OUT_F, G)
    orc.calibrate_threads(lambda: f(x, *layers[0]))
    for c, cb in layers:
        f(x, c, cb)
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        c, cb = layers[n % bufs]
        f(x, c, cb)
        n += 1
    dt = time.perf_counter() - t0
    cores = orc.num_threads()
    return {"value": n * algorithmic_bytes(IN_F, OUT_F) / dt / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
            "us_per_gemv": dt / n * 1e6,
            "sample": f"{n} GEMVs of {IN_F}x{OUT_F} over {bufs} distinct code matrices in {dt:.1f} s, {cores} OpenMP threads "
                      "(oracle/lutvq_oracle.c orc_aqlm_gemv_rowmajor = numba_gemv_lut_ restated, race-free over rows)"}


# ------------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--force-shard", action="store_true", help="experiment: take the sharded (push) path even at N=1")
    ap.add_argument("--tuning", default="", help="nw,pf,ctas_per_sm override for the GEMV kernel (experiments)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import rvv_lut_vq_b200 as pkg
    from rvv_lut_vq_b200 import ops, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = pkg.lib()
    if args.tuning:
        nw, pf, cps = [int(v) for v in args.tuning.split(",")]
        lib.lutvq_debug_set_gemv_tuning(nw, pf, cps)

    # ---- synthetic layers: NBUF distinct code matrices (per rank: its own output-channel shard) ----
    layers = [synth.make_layer_torch(IN_F, OUT_F, dev, seed=1000 * rank + i, scales="ones") for i in range(NBUF)]
    xs = torch.randn((NBUF, 1, IN_F), device=dev, dtype=torch.float16)
    ys = torch.empty((NBUF, 1, OUT_F), device=dev, dtype=torch.float16)
    stream = torch.cuda.Stream(dev)

    import ctypes
    P = lambda t: ctypes.c_void_p(t.data_ptr())

    # prepacked code images (what QuantizedLinear builds on first use) and a ring of workspaces: the NBUF GEMVs of a
    # step are independent (resident inputs), so they are issued with LUTVQ_FLAG_X_READY and overlap on the GPU
    packed = [ops.code2x8_prepack(L["codes"]) for L in layers]
    ws_need = (int(lib.lutvq_packed_workspace_bytes(OUT_F, IN_F)) + 255) // 256 * 256
    ws_ring = torch.zeros(ops.XREADY_WS_RING * ws_need, dtype=torch.uint8, device=dev)
    assert NBUF % ops.XREADY_WS_RING == 0
    FLAGS = {"x_ready": pkg._lib.FLAG_X_READY, "dependent": 0}

    # N > 1: each rank owns an OUT_F-row shard of an (N * OUT_F)-row layer; the GEMV kernel itself pushes its rows into
    # every peer's gather slot over NVLink (lutvq_sharded_matmat_packed) and one wait kernel per group of GROUP GEMVs
    # confirms that all N slices of those slots have landed locally.
    GROUP = int(os.environ.get("LUTVQ_BENCH_GROUP", "24"))      # 24 = one exchange per step (N=2: 3574 / 3791 / 4472 GB/s at 6 / 12 / 24)
    group = None
    sharded = world > 1 or args.force_shard
    if sharded and NBUF > 24:
        raise SystemExit("the sharded path has 24 gather slots: run --force-shard with LUTVQ_BENCH_NBUF=24")
    if sharded:
        from rvv_lut_vq_b200.sharded import ShardGroup
        group = ShardGroup.from_process_group(world * OUT_F, n_slots=NBUF) if world > 1 else ShardGroup(1, 0, OUT_F, n_slots=NBUF)
        slot_views = [group.slot_tensor(sl, world * OUT_F, torch.float16) for sl in range(NBUF)]

    def step_body(mode="x_ready", first=0, count=NBUF):
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        mask = 0
        for i in range(first, first + count):
            L = layers[i]
            ws = ws_ring[(i % ops.XREADY_WS_RING) * ws_need:]
            if not sharded:
                pkg._lib.check(lib.lutvq_code2x8_matmat_packed(P(xs[i]), P(packed[i]), P(L["codebooks"]), P(L["scales"]), None,
                                                               P(ys[i]), 1, OUT_F, IN_F, 0, FLAGS[mode], P(ws), ws_need, st))
            else:
                slot = i                                            # one gather slot per GEMV of the step: reused only after the step-end wait
                pkg._lib.check(lib.lutvq_sharded_matmat_packed(group._h, slot, rank * OUT_F, P(xs[i]), P(packed[i]), P(L["codebooks"]),
                                                               P(L["scales"]), None, OUT_F, IN_F, 0, FLAGS[mode], P(ws), ws_need, st))
                mask |= 1 << slot
                if i % GROUP == GROUP - 1:
                    # the exchange of the group runs on a side stream (a parallel branch of the captured graph), so the
                    # GEMV launch chain on the main stream is never queued behind it
                    main = torch.cuda.current_stream(dev)
                    side.wait_stream(main)
                    with torch.cuda.stream(side):
                        pkg._lib.check(lib.lutvq_shard_wait_mask(group._h, mask, ctypes.c_void_p(side.cuda_stream)))
                    mask = 0
        if sharded:
            torch.cuda.current_stream(dev).wait_stream(side)            # end of the (partial) step: every slice confirmed

    side = torch.cuda.Stream(dev) if sharded else None

    with torch.cuda.stream(stream):
        step_body()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            step_body()

        def run_step():
            graph.replay()

        def timed(n_steps):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(n_steps):
                run_step()
            e1.record(stream)
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            if world > 1:
                t = torch.tensor([ms], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            return ms

        for _ in range(args.warmup):
            run_step()
        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()
        ms_total = timed(args.steps)
        # keep the GPU under the same load a little longer so the 100 ms sampler sees it
        t_end = time.time() + 0.5
        while rank == 0 and world == 1 and time.time() < t_end:
            run_step()
        torch.cuda.synchronize()
        clocks = sampler.stop() if rank == 0 else None
        ms_per_step = ms_total / args.steps
        bytes_per_gemv = algorithmic_bytes(IN_F, OUT_F)
        # the same ring issued WITHOUT the X_READY promise: every GEMV waits for the previous one (decode-like chain)
        dep_us = None
        if world == 1:
            gdep = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gdep, stream=stream):
                step_body("dependent")
            for _ in range(5):
                gdep.replay()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = max(20, min(args.steps, 200))
            a.record(stream)
            for _ in range(reps):
                gdep.replay()
            b.record(stream)
            torch.cuda.synchronize()
            dep_us = a.elapsed_time(b) * 1e3 / (reps * NBUF)
        value = world * NBUF * bytes_per_gemv / (ms_per_step * 1e-3) / 1e9
        us_per_gemv = ms_per_step * 1e3 / NBUF

        # ---- e2e: host buffers through the reference-facing op ----
        x_host = torch.randn((NBUF, 1, IN_F), dtype=torch.float16).pin_memory()
        y_host = torch.empty((NBUF, 1, OUT_F), dtype=torch.float16).pin_memory()
        yg_host = torch.empty((NBUF, world * OUT_F), dtype=torch.float16).pin_memory() if sharded else None

        def e2e_step():
            # one step end to end: the step's activations come from pinned host memory, the C-ABI GEMVs run (as the
            # captured graph for 1 GPU, eagerly + all-gather otherwise), the step's outputs are read back to the host
            xs.copy_(x_host, non_blocking=True)
            run_step()
            if not sharded:
                y_host.copy_(ys, non_blocking=True)
            else:
                for j in range(NBUF):                                   # the gathered outputs live in the slot ring
                    yg_host[j].copy_(slot_views[j], non_blocking=True)

        e2e_steps = max(3, args.steps // 10)
        for _ in range(3):
            e2e_step()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(e2e_steps):
            e2e_step()
        e1.record(stream)
        torch.cuda.synchronize()
        e2e_ms = e0.elapsed_time(e1) / e2e_steps
        if world > 1:
            t = torch.tensor([e2e_ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_ms = float(t.item())
        e2e_val = world * NBUF * bytes_per_gemv / (e2e_ms * 1e-3) / 1e9
        # per-call cost of the torch-facing op (QuantizedLinear's path): H2D of one row, one op call, D2H of one row
        xd1 = x_host[0].to(dev)
        for _ in range(5):
            ops.code2x8_matmat_packed(xd1, packed[0], OUT_F, layers[0]["codebooks"], layers[0]["scales"], None)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n_calls = 200
        for i in range(n_calls):
            xd = x_host[i % NBUF].to(dev, non_blocking=True)
            yd = ops.code2x8_matmat_packed(xd, packed[i % NBUF], OUT_F, layers[i % NBUF]["codebooks"], layers[i % NBUF]["scales"], None)
            y_host[i % NBUF].copy_(yd, non_blocking=True)
        torch.cuda.synchronize()
        eager_call_us = (time.perf_counter() - t0) / n_calls * 1e6

        # ---- the other shapes of the configs[1] sweep (not part of the timed region above) ----
        sweep = []
        if rank == 0 and world == 1 and not args.no_sweep:
            peak, _ = measured_peak()
            for (i_f, o_f) in synth.SWEEP_7B:
                nb = -(-max(8, int(280e6 // (o_f * i_f // 4))) // ops.XREADY_WS_RING) * ops.XREADY_WS_RING
                Ls = [synth.make_layer_torch(i_f, o_f, dev, seed=77 + k, scales="ones") for k in range(nb)]
                xv = torch.randn((nb, 1, i_f), device=dev, dtype=torch.float16)
                yv = torch.empty((nb, 1, o_f), device=dev, dtype=torch.float16)

                pk = [ops.code2x8_prepack(L["codes"]) for L in Ls]
                need = (int(lib.lutvq_packed_workspace_bytes(o_f, i_f)) + 255) // 256 * 256
                ring = torch.zeros(ops.XREADY_WS_RING * need, dtype=torch.uint8, device=dev)

                def body(flags):
                    st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
                    for k, L in enumerate(Ls):
                        pkg._lib.check(lib.lutvq_code2x8_matmat_packed(P(xv[k]), P(pk[k]), P(L["codebooks"]), P(L["scales"]), None,
                                                                       P(yv[k]), 1, o_f, i_f, 0, flags, P(ring[(k % ops.XREADY_WS_RING) * need:]), need, st))
                row = {"shape": f"{i_f}->{o_f}", "layers_in_ring": nb}
                for name, flags in (("x_ready", pkg._lib.FLAG_X_READY), ("dependent", 0)):
                    body(flags)
                    torch.cuda.synchronize()
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=stream):
                        body(flags)
                    for _ in range(5):
                        g.replay()
                    torch.cuda.synchronize()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    reps = 50
                    a.record(stream)
                    for _ in range(reps):
                        g.replay()
                    b.record(stream)
                    torch.cuda.synchronize()
                    us = a.elapsed_time(b) * 1e3 / (reps * nb)
                    gbps = algorithmic_bytes(i_f, o_f) / (us * 1e-6) / 1e9
                    row[name] = {"us": round(us, 3), "gbps": round(gbps, 1), "frac_of_measured_peak": round(gbps / peak, 4)}
                    del g
                sweep.append(row)
                del Ls, pk

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    achieved = bytes_per_gemv / (us_per_gemv * 1e-6) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("lut_gemv_2x8_kernel_4096x11008_dram_bytes_per_launch")
    except Exception:
        pass
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": f"AQLM 2x8 LUT-GEMV {IN_F}x{OUT_F} batch 1 (Llama-2-7B gate_proj, BASELINE configs[1] headline shape)",
                   "gemvs_per_step": NBUF, "cache": f"inputs larger than L2: {NBUF} distinct code matrices = {NBUF * OUT_F * IN_F // 4 / 1e6:.0f} MB per step",
                   "launch": ("one CUDA graph per step; the GEMVs of a step are independent (resident inputs), issued with LUTVQ_FLAG_X_READY on "
                              "prepacked codes + a ring of 8 workspaces, so consecutive launches overlap under programmatic dependent launch; "
                              "dependent_chain_us_per_gemv is the same ring without the promise") if world == 1 else ("one CUDA graph per step; every GEMV (lutvq_sharded_matmat_packed, X_READY ring) stores its rows into all peers' gather slots over NVLink "
                              "inside the kernel; per group of 6 GEMVs one exchange launch on a side stream (parallel graph branch) forwards the finished slices to the peers and confirms the group's all-gather locally (lutvq_shard_wait_mask); the step ends with every slice confirmed; no NCCL on the data path"),
                   "sharding": "none" if world == 1 else f"output channels, {world} x {OUT_F} rows"},
        "us_per_gemv": us_per_gemv,
        "dependent_chain_us_per_gemv": dep_us,
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": NBUF * IN_F * 2, "d2h_bytes_per_step": NBUF * world * OUT_F * 2,
                "us_per_gemv": e2e_ms * 1e3 / NBUF, "eager_torch_op_call_us": eager_call_us,
                "how": "per step: one H2D copy of the step's NBUF activation rows from pinned host memory, the step's C-ABI GEMVs, one D2H copy of the NBUF output rows"},
        "gpu_launches": args.steps * (NBUF + (NBUF // GROUP if world > 1 else 0)),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                     "kernel": "lut_gemv_2x8_kernel", "algorithmic_bytes_per_launch": bytes_per_gemv, "peak_source": peak_src,
                     "frac_of_nominal_8TBs": achieved / 8000.0},
        "clocks": clocks,
        "sweep": sweep,
    }
    if not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline()
    elif world == 1:
        out["cpu_baseline"] = None
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
