#!/usr/bin/env python
"""bench.py -- LUT-VQ GEMV throughput + Llama decode tok/s on B200 (BASELINE.json metric: "LUT-VQ GEMV us & HBM GB/s (AQLM 2x8,
4096x11008); Llama-2-7B decode tok/s"), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE configs[1] headline shape): Llama-2-7B gate_proj 4096->11008, AQLM 2x8 codes, batch 1.
One STEP = one LUT-GEMV over each of 64 distinct layers per GPU (random codes/codebooks/scales = the gate/up projections of one
Llama-2-7B token; 721 MB of codes per GPU and step, > 2x the 126 MB L2, so every launch streams its codes from HBM), replayed as
ONE CUDA graph -- the same 64 launches per graph at every N.
  value     whole-job GB/s on the algorithmic byte stream (O*(I/8)*2 + 8192 + 2I + 4O per GEMV), inputs resident in
            HBM, timed with CUDA events on the launching stream, max over ranks;
  e2e       the same step with HOST buffers: one H2D copy of the step's 64 activation rows from pinned memory, the step's
            C-ABI GEMVs, one D2H copy of the 64 output rows, all inside the timed region;
  dependent_chain_us_per_gemv   the same ring WITHOUT the X_READY promise: what a decoder issues;
  decode    Llama-2-7B (N=1) and Llama-2-70B (every N: strong scaling, codes sharded by output channel) decode tok/s through
            decode.Decoder, 128 new tokens, batch 1 (benchmark/AQLM/benchmark/generate_benchmark.py:95-106);
  prefill / dequant / kernel_to_beat (N=1)   configs[2] batch 16..256, Code2x8Dequant, and the reference's own CUDA kernels
            (baseline/_ref/aqlm_cuda, built unmodified by tools/vendor_ref_cuda.py) timed on the same GPU;
  parity    N>1: before timing, every rank's gathered slot is compared bit for bit with the unsharded GEMV of the concatenated
            layer; a mismatch exits non-zero;
  roofline / cpu_baseline / clocks / sweep: see DESIGN.md "Measurement".
N>1 (torchrun, one rank per GPU): weak scaling of the GEMV ring -- each rank owns an 11008-row shard of an N*11008-row layer and
every output slice is all-gathered into every rank's gather slot by the push exchange over NVLink (no NCCL on the data path).
--impl reference: the reference's CPU LUT-GEMV (oracle port of numba_gemv_lut_, all host threads) on the same config.
"""
from __future__ import annotations

import argparse
import ctypes
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
NBUF = int(os.environ.get("LUTVQ_BENCH_NBUF", "0")) or 64     # GEMVs per step and per GPU, at every N
METRIC = "lutvq_gemv_hbm_gbps"
UNIT = "GB/s"
WORKLOAD = f"AQLM 2x8 LUT-GEMV {IN_F}x{OUT_F} batch 1 (Llama-2-7B gate_proj, BASELINE configs[1] headline shape), {NBUF} distinct layers per step"
DECODE_TOKENS = 128                 # generate_benchmark.py:95-106: 128 new tokens, batch 1


def algorithmic_bytes(i: int, o: int) -> int:
    return o * (i // 8) * 2 + 2 * 256 * 8 * 2 + 2 * i + 4 * o


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops", 1650.0)), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, 1650.0, "fallback (B200_PROFILING.md)"


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


# ------------------------------------------------------------------------------------------------ reference arm / CPU baseline
def _cpu_layers(n):
    rng = np.random.default_rng(0)
    G = IN_F // 8
    layers = [(rng.integers(0, 256, size=(OUT_F, G, 2), dtype=np.uint8).view(np.int8),
               rng.standard_normal((2, 256, 8)).astype(np.float16).view(np.uint16),
               np.ones(OUT_F, dtype=np.float16).view(np.uint16)) for _ in range(n)]
    x = rng.standard_normal(IN_F).astype(np.float16).view(np.uint16)
    return layers, x


def run_reference(args):
    """The reference's CPU implementation of the path (numba_gemv_lut_ semantics, aqlm numba_kernel.py:38-48), timed as the
    oracle port on all host threads.  Same workload as our arm: one step = NBUF GEMVs of the headline shape; the host keeps
    CPU_BUFS distinct code matrices (> the CPU's last-level cache) and walks them round-robin -- a bounded sample of the step's
    NBUF matrices, stated in cpu_baseline.sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as orc
    CPU_BUFS = 16
    layers, x = _cpu_layers(CPU_BUFS)
    f = orc.RowMajorGemv(OUT_F, IN_F // 8)
    orc.calibrate_threads(lambda: f(x, *layers[0]))
    step = lambda: [f(x, *layers[i % CPU_BUFS]) for i in range(NBUF)]
    for _ in range(min(args.warmup, 3)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    ms = dt / args.steps * 1e3
    val = NBUF * algorithmic_bytes(IN_F, OUT_F) / (ms * 1e-3) / 1e9
    cores = orc.num_threads()
    sample = (f"{args.steps} steps x {NBUF} GEMVs of {IN_F}x{OUT_F} walking {CPU_BUFS} distinct code matrices (180 MB) round-robin, "
              f"{cores} OpenMP threads (oracle/lutvq_oracle.c orc_aqlm_gemv_rowmajor = numba_gemv_lut_ restated, race-free over rows)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "gemvs_per_step": NBUF},
        "arm": {"path": "CPU LUT-GEMV (oracle port of the reference's numba kernel)", "distinct_code_matrices_resident": CPU_BUFS, "warmup_steps_run": min(args.warmup, 3)},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "us_per_gemv": ms * 1e3 / NBUF,
    }))


def cpu_baseline(seconds: float = 12.0):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as orc
    bufs = 16                                     # the same 180 MB working set as the reference arm (larger than the host's last-level cache)
    layers, x = _cpu_layers(bufs)
    f = orc.RowMajorGemv(OUT_F, IN_F // 8)
    orc.calibrate_threads(lambda: f(x, *layers[0]))
    for L in layers:
        f(x, *L)
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds:
        f(x, *layers[n % bufs])
        n += 1
    dt = time.perf_counter() - t0
    cores = orc.num_threads()
    return {"value": n * algorithmic_bytes(IN_F, OUT_F) / dt / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
            "us_per_gemv": dt / n * 1e6,
            "sample": f"{n} GEMVs of {IN_F}x{OUT_F} over {bufs} distinct code matrices in {dt:.1f} s, {cores} OpenMP threads "
                      "(oracle/lutvq_oracle.c orc_aqlm_gemv_rowmajor = numba_gemv_lut_ restated, race-free over rows)"}


# ------------------------------------------------------------------------------------------------ secondary measurements
def graph_time_us(torch, stream, body, reps, inner=1):
    """us per `inner` unit: body() captured once as a CUDA graph on `stream`, replayed reps times between two events"""
    with torch.cuda.stream(stream):
        body()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            body()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            g.replay()
        b.record(stream)
        torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / (reps * inner)


def measure_decode(torch, dist, model: str, world: int, rank: int, dev, tokens: int = DECODE_TOKENS, exchange: str = "ll"):
    """decode tok/s of the Llama-shaped AQLM 2x8 decoder (decode.py): one CUDA graph per token, CUDA events, max over ranks"""
    from rvv_lut_vq_b200.decode import LLAMA2_7B, LLAMA2_70B, Decoder
    from rvv_lut_vq_b200.sharded import ShardGroup
    cfg = {"7b": LLAMA2_7B, "70b": LLAMA2_70B}[model]
    if world == 1:
        exchange = "push"                                        # one GPU: the plain decoder (no gather slots at all)
    group = ShardGroup.from_process_group(max(cfg.intermediate, cfg.hidden), n_slots=8) if world > 1 else None
    dec = Decoder(cfg, dev, world, rank, group, max_seq=tokens + 16, exchange=exchange)
    st = torch.cuda.Stream(dev)
    dec.reset(1)
    dec.capture(st)
    with torch.cuda.stream(st):
        for _ in range(3):
            dec.run_token()
        dec.reset(1)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(tokens):
            dec.run_token()
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    err = group.error() if group is not None else 0
    last_token = int(dec.token.item())
    us_tok = ms * 1e3 / tokens
    gemv_bytes = cfg.gemv_bytes_per_token()
    peak, _, _ = measured_peaks()
    out = {"model": cfg.name, "tok_per_s": 1e6 / us_tok, "us_per_token": us_tok, "tokens": tokens, "batch": 1, "n_gpus": world,
           "launches_per_token": dec.launches_per_token, "graph": dec.graph is not None,
           "gemv_code_stream_bytes_per_token": gemv_bytes,
           "frac_of_hbm_peak": gemv_bytes / world / (us_tok * 1e-6) / 1e9 / peak,
           "exchange": None if world == 1 else
                       {"ll": "LL words: every finished row is stored as {value | epoch} into all ranks' gather slots over NVLink by the GEMV itself; "
                              "the consuming kernels poll the words they read (no exchange launches, no flags, no fences); 4 exchanges per layer",
                        "push": "push: one exchange launch per GEMV group forwards the slices over NVLink and waits on epoch flags; 4 per layer",
                        "nccl": "nccl: all_gather_into_tensor per linear (eager)"}[exchange],
           "wait_errors": err, "last_token": last_token}
    del dec, group
    torch.cuda.empty_cache()
    return out


def measure_prefill_dequant(torch, dev, bf16_peak, hbm_peak):
    """BASELINE configs[2]: dequantize-to-smem + tcgen05 GEMM at batch 16..256 on 4096x11008, and the stand-alone dequantize"""
    from rvv_lut_vq_b200 import ops, synth
    st = torch.cuda.Stream(dev)
    NB = 24
    Ls = [synth.make_layer_torch(IN_F, OUT_F, dev, seed=900 + k, scales="ones") for k in range(NB)]
    rows = []
    for batch in (16, 32, 64, 128, 256, 2048):
        x = torch.randn((batch, IN_F), device=dev, dtype=torch.float16)

        def body_of(fn):
            def body():
                for L in Ls:
                    fn(x, L["codes"], L["codebooks"], L["scales"], None)
            return body
        # the op a user calls (fused tcgen05 kernel below ops.TWO_STEP_MIN_ROWS rows, dequantize-once + library GEMM above) and both routes
        us = graph_time_us(torch, st, body_of(ops.code2x8_matmat_dequant), 5, NB)
        us_fused = graph_time_us(torch, st, body_of(ops.code2x8_matmat_dequant_fused), 5, NB)
        us_two = graph_time_us(torch, st, body_of(ops.code2x8_matmat_dequant_two_step), 5, NB)
        flops = 2.0 * batch * IN_F * OUT_F
        rows.append({"batch": batch, "us": round(us, 2), "route": "two_step" if batch >= ops.two_step_min_rows(OUT_F, IN_F) else "fused",
                     "fused_tcgen05_us": round(us_fused, 2), "dequant_plus_library_gemm_us": round(us_two, 2),
                     "tflops": round(flops / us / 1e6, 1), "frac_of_bf16_peak": round(flops / us / 1e6 / bf16_peak, 4),
                     "code_stream_GBps": round(OUT_F * IN_F / 4 / us / 1e3, 1)})
    W = [torch.empty((OUT_F, IN_F), dtype=torch.float16, device=dev) for _ in range(2)]
    lib = __import__("rvv_lut_vq_b200").lib()
    P = lambda t: ctypes.c_void_p(t.data_ptr())

    def dq():
        s = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        for k, L in enumerate(Ls):
            lib.lutvq_code2x8_dequant(P(L["codes"]), P(W[k % 2]), P(L["codebooks"]), P(L["scales"]), OUT_F, IN_F, 0, s)
    us = graph_time_us(torch, st, dq, 5, NB)
    nbytes = OUT_F * IN_F // 4 + 2 * OUT_F * IN_F + 8192 + 2 * OUT_F
    dq_row = {"shape": f"{IN_F}x{OUT_F}", "us": round(us, 2), "GBps": round(nbytes / us / 1e3, 1), "frac_of_hbm_peak": round(nbytes / us / 1e3 / hbm_peak, 4),
              "algorithmic_bytes": nbytes}
    del Ls, W
    torch.cuda.empty_cache()
    return rows, dq_row


def measure_small_batch(torch, dev):
    """batch 1..8 through lutvq_code2x8_matmat_packed on 4096x11008 (a dependent chain over a ring of L2-cold layers): us per call with
    the rows of a call grouped into one launch (what ops / QuantizedLinear do) and with one launch per row (the reference's loop)"""
    from rvv_lut_vq_b200 import ops, synth
    lib = __import__("rvv_lut_vq_b200").lib()
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    st = torch.cuda.Stream(dev)
    NB = 32
    Ls = [synth.make_layer_torch(IN_F, OUT_F, dev, seed=700 + k, scales="randn") for k in range(NB)]
    pk = [ops.code2x8_prepack(L["codes"]) for L in Ls]
    need = (int(lib.lutvq_packed_workspace_bytes(OUT_F, IN_F)) + 255) // 256 * 256
    ws = torch.zeros(need * 8, dtype=torch.uint8, device=dev)
    rows = []
    for batch in (1, 2, 4, 6, 8):
        x = torch.randn((batch, IN_F), device=dev, dtype=torch.float16)
        y = torch.empty((batch, OUT_F), device=dev, dtype=torch.float16)
        r = {"batch": batch}
        for name, nbytes in (("us", need * min(batch, 8)), ("one_launch_per_row_us", need)):
            def body():
                s = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
                for k, L in enumerate(Ls):
                    rc = lib.lutvq_code2x8_matmat_packed(P(x), P(pk[k]), P(L["codebooks"]), P(L["scales"]), None, P(y), batch, OUT_F, IN_F, 0, 0,
                                                         P(ws), nbytes, s)
                    assert rc == 0, lib.lutvq_last_error()
            r[name] = round(graph_time_us(torch, st, body, 5, NB), 2)
        rows.append(r)
    for r in rows:
        r["vs_batch1"] = round(r["us"] / rows[0]["us"], 2)
    del Ls, pk, ws
    torch.cuda.empty_cache()
    return rows


def kernel_to_beat():
    """the reference's own CUDA kernels on this GPU (baseline/_ref/aqlm_cuda, tools/ref_cuda_bench.py); never fatal"""
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import contextlib
        import ref_cuda_bench
        with contextlib.redirect_stdout(sys.stderr):                 # stdout carries exactly one JSON line
            return ref_cuda_bench.run(quick=True, batches=(16, 256))
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"{type(e).__name__}: {e}"[:300]}


# ------------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sweep", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip decode / prefill / dequant / kernel_to_beat (kernel experiments)")
    ap.add_argument("--force-shard", action="store_true", help="experiment: take the sharded (push) path even at N=1")
    ap.add_argument("--tuning", default="", help="nw,nslot override for the GEMV kernel (experiments)")
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
        t = [int(v) for v in args.tuning.split(",")] + [0, 0]
        lib.lutvq_debug_set_gemv_tuning(t[0], t[1], t[2])
    hbm_peak, bf16_peak, peak_src = measured_peaks()

    # ---- synthetic layers: NBUF distinct code matrices (per rank: its own output-channel shard) ----
    layers = [synth.make_layer_torch(IN_F, OUT_F, dev, seed=1000 * rank + i, scales="ones") for i in range(NBUF)]
    xs = torch.randn((NBUF, 1, IN_F), device=dev, dtype=torch.float16)
    if world > 1:
        dist.broadcast(xs, 0)                                     # activations and codebooks are replicated; codes are per rank
        for L in layers:
            dist.broadcast(L["codebooks"], 0)
    ys = torch.empty((NBUF, 1, OUT_F), device=dev, dtype=torch.float16)
    stream = torch.cuda.Stream(dev)
    P = lambda t: ctypes.c_void_p(t.data_ptr())

    # prepacked code images (what QuantizedLinear builds on first use) and a ring of workspaces: the NBUF GEMVs of a
    # step are independent (resident inputs), so they are issued with LUTVQ_FLAG_X_READY and overlap on the GPU
    packed = [ops.code2x8_prepack(L["codes"]) for L in layers]
    ws_need = (int(lib.lutvq_packed_workspace_bytes(OUT_F, IN_F)) + 255) // 256 * 256
    ws_ring = torch.zeros(ops.XREADY_WS_RING * ws_need, dtype=torch.uint8, device=dev)
    assert NBUF % ops.XREADY_WS_RING == 0 and NBUF <= 64
    FLAGS = {"x_ready": pkg._lib.FLAG_X_READY, "dependent": 0}

    # N > 1: each rank owns an OUT_F-row shard of an (N * OUT_F)-row layer; every GEMV fills the rank's slice of its own
    # gather slot, and ONE exchange launch per step (side stream = parallel graph branch) forwards all slices to the peers
    # over NVLink and confirms that all N slices of every slot have landed locally.
    GROUP = int(os.environ.get("LUTVQ_BENCH_GROUP", str(NBUF)))
    group = None
    sharded = world > 1 or args.force_shard
    parity = None
    if sharded:
        from rvv_lut_vq_b200.sharded import ShardGroup
        group = ShardGroup.from_process_group(world * OUT_F, n_slots=NBUF) if world > 1 else ShardGroup(1, 0, OUT_F, n_slots=NBUF)
        slot_views = [group.slot_tensor(sl, world * OUT_F, torch.float16) for sl in range(NBUF)]
    side = torch.cuda.Stream(dev) if sharded else None

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
                    main_s = torch.cuda.current_stream(dev)
                    side.wait_stream(main_s)
                    with torch.cuda.stream(side):
                        pkg._lib.check(lib.lutvq_shard_wait_mask(group._h, mask, ctypes.c_void_p(side.cuda_stream)))
                    mask = 0
        if sharded:
            torch.cuda.current_stream(dev).wait_stream(side)            # end of the (partial) step: every slice confirmed

    with torch.cuda.stream(stream):
        step_body()
        torch.cuda.synchronize()
        if world > 1:
            # ---- parity of the multi-GPU data plane, before any timing: every gathered slot == the unsharded GEMV of the
            # concatenated layer (same kernel, prob_m = N * OUT_F), bit for bit, on every rank
            dist.barrier()
            n_check = min(NBUF, int(os.environ.get("LUTVQ_BENCH_PARITY_SLOTS", "8")))
            ok = True
            for i in range(n_check):
                parts = [torch.empty_like(layers[i]["codes"]) for _ in range(world)]
                dist.all_gather(parts, layers[i]["codes"])
                whole = torch.cat(parts)                                        # the concatenated layer: [N * OUT_F, G, 2]
                want = ops.code2x8_matmat_packed(xs[i], ops.code2x8_prepack(whole), world * OUT_F, layers[i]["codebooks"],
                                                 torch.ones(world * OUT_F, dtype=torch.float16, device=dev), None).reshape(-1)
                ok = ok and bool(torch.equal(want.view(torch.int16), slot_views[i].view(torch.int16)))
                del parts, whole
            flag = torch.tensor([1 if ok else 0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            parity = {"gather_bit_identical": bool(flag.item() == 1), "slots_checked": n_check, "ranks": world,
                      "how": "each rank's gathered slot vs lutvq_code2x8_matmat_packed on the concatenated layer (codes all-gathered by NCCL for the check only), bit for bit"}
            if not parity["gather_bit_identical"]:
                if rank == 0:
                    print(json.dumps({"metric": METRIC, "parity": parity, "error": "gathered slots differ from the unsharded result"}))
                dist.destroy_process_group()
                sys.exit(3)
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

        for _ in range(max(args.warmup, 3)):
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
            dep_us = graph_time_us(torch, stream, lambda: step_body("dependent"), max(20, min(args.steps, 100)), NBUF)
        value = world * NBUF * bytes_per_gemv / (ms_per_step * 1e-3) / 1e9
        us_per_gemv = ms_per_step * 1e3 / NBUF

        # ---- e2e: host buffers ----
        x_host = torch.randn((NBUF, 1, IN_F), dtype=torch.float16).pin_memory()
        y_host = torch.empty((NBUF, 1, OUT_F), dtype=torch.float16).pin_memory()
        yg_host = torch.empty((NBUF, world * OUT_F), dtype=torch.float16).pin_memory() if sharded else None

        def e2e_step():
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
        # the op call alone (input already on the device, no copies): host cost of QuantizedLinear's per-token call
        t0 = time.perf_counter()
        for i in range(n_calls):
            ops.code2x8_matmat_packed(xd1, packed[i % NBUF], OUT_F, layers[i % NBUF]["codebooks"], layers[i % NBUF]["scales"], None)
        torch.cuda.synchronize()
        eager_op_only_us = (time.perf_counter() - t0) / n_calls * 1e6

        # ---- the other shapes of the configs[1] sweep (not part of the timed region above) ----
        sweep = []
        if rank == 0 and world == 1 and not args.no_sweep:
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
                    us = graph_time_us(torch, stream, lambda: body(flags), 50, nb)
                    gbps = algorithmic_bytes(i_f, o_f) / (us * 1e-6) / 1e9
                    row[name] = {"us": round(us, 3), "gbps": round(gbps, 1), "frac_of_measured_peak": round(gbps / hbm_peak, 4)}
                sweep.append(row)
                del Ls, pk

    # free the ring before the model-sized measurements
    del layers, packed, graph
    torch.cuda.empty_cache()

    # ---- decode tok/s (the second half of BASELINE.json's metric) ----
    decode = decode70 = prefill = dequant = ktb = small_batch = None
    if not args.no_extras:
        if world == 1:
            decode = measure_decode(torch, dist, "7b", 1, 0, dev)
            decode70 = measure_decode(torch, dist, "70b", 1, 0, dev)
            prefill, dequant = measure_prefill_dequant(torch, dev, bf16_peak, hbm_peak)
            small_batch = measure_small_batch(torch, dev)
            ktb = kernel_to_beat()
        else:
            decode70 = measure_decode(torch, dist, "70b", world, rank, dev, exchange=os.environ.get("LUTVQ_BENCH_EXCHANGE", "ll"))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    achieved = bytes_per_gemv / (us_per_gemv * 1e-6) / 1e9
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        traffic = tj.get("lut_gemv_2x8_kernel_4096x11008_dram_bytes_per_launch")
        traffic_src = f"profiles/traffic.json (ncu --set full capture, {tj.get('when', 'see file')}); a citation, not measured in this run"
    except Exception:
        pass
    n_exch = (NBUF // GROUP) if sharded else 0
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": WORKLOAD, "gemvs_per_step": NBUF,
                   "cache": f"inputs larger than L2: {NBUF} distinct code matrices = {NBUF * OUT_F * IN_F // 4 / 1e6:.0f} MB per GPU and step",
                   "launch": ("one CUDA graph per step; the GEMVs of a step are independent (resident inputs), issued with LUTVQ_FLAG_X_READY on "
                              "prepacked codes + a ring of 8 workspaces, so consecutive launches overlap under programmatic dependent launch; "
                              "dependent_chain_us_per_gemv is the same ring without the promise") if not sharded else
                             (f"one CUDA graph per step of {NBUF} GEMVs per rank (lutvq_sharded_matmat_packed, X_READY ring): each fills the rank's slice of its "
                              f"own gather slot; {n_exch} exchange launch(es) per step on a side stream (parallel graph branch) forward the finished slices to "
                              "the peers with NVLink stores, publish the epochs and confirm all N slices of every slot locally (lutvq_shard_wait_mask); "
                              "no NCCL on the data path"),
                   "sharding": "none" if world == 1 else f"output channels, {world} x {OUT_F} rows"},
        "us_per_gemv": us_per_gemv,
        "dependent_chain_us_per_gemv": dep_us,
        "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": NBUF * IN_F * 2, "d2h_bytes_per_step": NBUF * world * OUT_F * 2,
                "us_per_gemv": e2e_ms * 1e3 / NBUF, "eager_torch_op_call_us": eager_call_us, "eager_torch_op_only_us": eager_op_only_us,
                "how": "per step: one H2D copy of the step's NBUF activation rows from pinned host memory, the step's C-ABI GEMVs, one D2H copy of the NBUF output rows"},
        "gpu_launches": args.steps * (NBUF + n_exch),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": traffic,
                     "traffic_source": traffic_src,
                     "kernel": "lut_gemv_2x8_kernel", "algorithmic_bytes_per_launch": bytes_per_gemv, "peak_source": peak_src,
                     "frac_of_nominal_8TBs": achieved / 8000.0},
        "clocks": clocks,
        "sweep": sweep,
    }
    if parity is not None:
        out["parity"] = parity
    if decode is not None:
        out["decode"] = decode
    if decode70 is not None:
        out["decode_70b"] = decode70
    if small_batch is not None:
        out["small_batch"] = {"shape": f"{IN_F}x{OUT_F}", "rows": small_batch}
    if prefill is not None:
        out["prefill"] = prefill
        out["dequant"] = dequant
    if ktb is not None:
        out["kernel_to_beat"] = ktb
    if not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline()
    elif world == 1:
        out["cpu_baseline"] = None
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
