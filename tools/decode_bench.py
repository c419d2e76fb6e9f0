#!/usr/bin/env python
"""Decode tok/s of the Llama-shaped AQLM 2x8 decoder (BASELINE.json configs[3], [4]).
    python tools/decode_bench.py --model 7b [--tokens 128]
    torchrun --nproc-per-node N ... tools/decode_bench.py --model 70b --exchange push|nccl
One JSON line per run on rank 0: tok/s (max over ranks, CUDA events), us per token, GEMV share of the HBM roofline."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="7b", choices=["7b", "70b", "tiny"])
    ap.add_argument("--tokens", type=int, default=128)
    ap.add_argument("--exchange", default="ll", choices=["ll", "push", "nccl"])
    ap.add_argument("--ungrouped", action="store_true", help="one launch per linear instead of grouped q/k/v and gate/up launches")
    ap.add_argument("--unfused-swiglu", action="store_true", help="separate silu*up launch instead of the fused down_proj activation load")
    ap.add_argument("--plain", action="store_true", help="one GPU: plain dependent launches instead of the LL hand-over")
    ap.add_argument("--dataflow", action="store_true", help="LL exchange with the GEMVs as dataflow launches (they poll their LL activations instead of waiting)")
    ap.add_argument("--tuning", default="", help="nw,pf,pa for every GEMV (lutvq_debug_set_gemv_tuning)")
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (memory-bounded experiments)")
    args = ap.parse_args()
    import torch.distributed as dist
    from rvv_lut_vq_b200.decode import Decoder, LLAMA2_7B, LLAMA2_70B, TINY
    from rvv_lut_vq_b200.sharded import ShardGroup
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cfg = {"7b": LLAMA2_7B, "70b": LLAMA2_70B, "tiny": TINY}[args.model]
    if args.layers:
        import dataclasses
        cfg = dataclasses.replace(cfg, n_layers=args.layers)
    if world > 1 and args.exchange in ("push", "ll"):
        group = ShardGroup.from_process_group(max(cfg.intermediate, cfg.hidden), n_slots=16)
    elif world == 1 and args.exchange == "ll" and not args.plain:
        group = ShardGroup(1, 0, max(cfg.intermediate, cfg.hidden), n_slots=16)       # one rank: the LL words are the kernel-to-kernel hand-over
    else:
        group = None
    if args.tuning:
        import rvv_lut_vq_b200 as r
        r.lib().lutvq_debug_set_gemv_tuning(*[int(v) for v in args.tuning.split(",")])
    exchange = args.exchange if group is not None or world > 1 else "push"
    dec = Decoder(cfg, dev, world, rank, group, max_seq=args.tokens + 16, exchange=exchange, grouped=not args.ungrouped, fuse_swiglu=not args.unfused_swiglu,
                  dataflow=args.dataflow)
    st = torch.cuda.Stream()
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
        for _ in range(args.tokens):
            dec.run_token()
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    err = group.error() if group is not None else 0
    if rank == 0:
        us_tok = ms * 1e3 / args.tokens
        gemv_bytes = cfg.gemv_bytes_per_token()
        peak = 6555.8
        try:
            peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
        except Exception:
            pass
        print(json.dumps({"metric": "decode_tok_per_s", "model": cfg.name, "layers": cfg.n_layers, "n_gpus": world, "exchange": exchange if group is not None or world > 1 else None, "dataflow": bool(dec.dataflow and exchange == "ll"), "tuning": args.tuning or None, "grouped_launches": not args.ungrouped, "fused_swiglu": dec.fuse_swiglu,
                          "value": 1e6 / us_tok, "unit": "tok/s", "us_per_token": us_tok, "tokens": args.tokens, "batch": 1,
                          "launches_per_token": dec.launches_per_token, "graph": dec.graph is not None,
                          "gemv_code_stream_bytes_per_token": gemv_bytes,
                          "gemv_stream_GBps_aggregate": gemv_bytes / (us_tok * 1e-6) / 1e9,
                          "frac_of_hbm_peak_per_gpu": gemv_bytes / world / (us_tok * 1e-6) / 1e9 / peak, "wait_errors": err}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
