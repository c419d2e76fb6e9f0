#!/usr/bin/env python
"""Where a decode token's time goes (Llama-2-7B shapes, one GPU): the full step against the same step with only its
GEMV launches / only its glue launches, each captured as one CUDA graph.  Timing only (partial chains compute nonsense)."""
import ctypes, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rvv_lut_vq_b200 import _lib
from rvv_lut_vq_b200.decode import Decoder, LLAMA2_7B

dev = torch.device("cuda", 0)
dec = Decoder(LLAMA2_7B, dev, max_seq=160)
lib, cfg = dec.lib, dec.cfg
P = lambda t: ctypes.c_void_p(t.data_ptr())
st = torch.cuda.Stream()
sp = lambda: ctypes.c_void_p(st.cuda_stream)
dec._st = sp
dec.reset(1)
dec.pos.fill_(96)


def chain(gemv=True, glue=True, which=("qkv", "o", "gu", "down"), parts=("rms1", "attn", "rms2")):
    delta = None
    for L in dec.layers:
        if glue and "rms1" in parts:
            _lib.check(lib.lutvq_decode_add_rmsnorm(P(dec.resid), None if delta is None else P(delta), P(L["norm1"]), P(dec.normed), cfg.hidden, cfg.eps, sp()))
        q, k, v = dec._group_gemv(L, ["q", "k", "v"], dec.normed) if gemv and "qkv" in which else (dec.bufs["q"], dec.bufs["k"], dec.bufs["v"])
        if glue and "attn" in parts:
            _lib.check(lib.lutvq_decode_attention(P(q), P(k), P(v), P(L["kcache"]), P(L["vcache"]), P(dec.pos), P(dec.attn_out), cfg.n_heads,
                                                  cfg.n_kv_heads, cfg.head_dim, dec.max_seq, cfg.rope_theta, sp()))
        (o,) = dec._group_gemv(L, ["o"], dec.attn_out) if gemv and "o" in which else (dec.bufs["o"],)
        if glue and "rms2" in parts:
            _lib.check(lib.lutvq_decode_add_rmsnorm(P(dec.resid), P(o), P(L["norm2"]), P(dec.normed), cfg.hidden, cfg.eps, sp()))
        gate, up = dec._group_gemv(L, ["gate", "up"], dec.normed) if gemv and "gu" in which else (dec.bufs["gate"], dec.bufs["up"])
        (delta,) = dec._group_gemv(L, ["down"], up, gate=gate) if gemv and "down" in which else (dec.bufs["down"],)


def timed(**kw):
    with torch.cuda.stream(st):
        chain(**kw)
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            chain(**kw)
        for _ in range(3):
            g.replay()
        st.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for _ in range(20):
            g.replay()
        b.record(st)
        st.synchronize()
    return a.elapsed_time(b) * 1e3 / 20 / cfg.n_layers


out = {"model": cfg.name, "pos": 96, "unit": "us per layer",
       "full": timed(), "gemv_only": timed(glue=False), "glue_only": timed(gemv=False)}
for w in ("qkv", "o", "gu", "down"):
    out["gemv_only_" + w] = timed(glue=False, which=(w,))
for pp in ("rms1", "attn", "rms2"):
    out["full_without_" + pp] = timed(parts=tuple(x for x in ("rms1", "attn", "rms2") if x != pp))
for name, t in (("all_3cta", (8, 1, 1)), ("all_2cta", (8, 2, 2))):
    lib.lutvq_debug_set_gemv_tuning(*t)
    out["gemv_only_" + name] = timed(glue=False)
    out["full_" + name] = timed()
lib.lutvq_debug_set_gemv_tuning(0, 0, 0)
print(json.dumps({k: (round(v, 2) if isinstance(v, float) else v) for k, v in out.items()}))
