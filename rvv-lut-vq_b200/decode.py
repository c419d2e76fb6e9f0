"""Llama-shaped batch-1 decoder over AQLM 2x8 linears (BASELINE.json configs[3]: Llama-2-7B, configs[4]: Llama-2-70B
sharded by output channel).  Random codes / codebooks / scales per layer, DISTINCT buffers for every layer (the
reference's generate_benchmark.py:67-79 aliases one layer, which would sit in L2), fp16 lm_head as in the AQLM checkpoints.

One token = one CUDA graph replay: per layer  add+RMSNorm -> q,k,v GEMV -> RoPE/KV-append/attention -> o GEMV ->
add+RMSNorm -> gate,up GEMV -> SwiGLU -> down GEMV; then final norm, lm_head (cuBLAS GEMV, plain library call), argmax.
Every linear goes through liblutvq's C-ABI: ``lutvq_code2x8_matmat_packed`` on one GPU, ``lutvq_sharded_matmat_packed``
(+ ``lutvq_shard_wait``) when sharded -- each rank then holds 1/P of every linear's rows and the whole activations.
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass
from typing import List, Optional

import torch

from . import _lib, ops, synth
from .sharded import ShardGroup, shard_rows


@dataclass
class DecoderConfig:
    name: str
    hidden: int
    intermediate: int
    n_layers: int
    n_heads: int
    n_kv_heads: int
    vocab: int = 32000
    head_dim: int = 128
    rope_theta: float = 10000.0
    eps: float = 1e-5

    def linear_shapes(self):
        kv = self.n_kv_heads * self.head_dim
        return {"q": (self.hidden, self.hidden), "k": (self.hidden, kv), "v": (self.hidden, kv), "o": (self.hidden, self.hidden),
                "gate": (self.hidden, self.intermediate), "up": (self.hidden, self.intermediate), "down": (self.intermediate, self.hidden)}

    def gemv_bytes_per_token(self) -> int:
        return self.n_layers * sum(synth.gemv_algorithmic_bytes(i, o) for (i, o) in self.linear_shapes().values())


LLAMA2_7B = DecoderConfig("llama-2-7b", 4096, 11008, 32, 32, 32)
LLAMA2_70B = DecoderConfig("llama-2-70b", 8192, 28672, 80, 64, 8)
TINY = DecoderConfig("tiny", 512, 1024, 2, 4, 2, vocab=512)


class _Linear:
    """one AQLM 2x8 linear: this rank's rows (packed), shared codebooks; writes its output into ``out`` (the caller's
    buffer on one GPU, a gather slot when sharded)"""

    def __init__(self, in_f: int, out_f: int, dev, seed: int, world: int, rank: int, keep_codes: bool = False, skeleton: bool = False):
        self.in_f, self.out_f = in_f, out_f
        self.begin, self.end = shard_rows(out_f, world, rank)
        self.rows = self.end - self.begin
        if skeleton:                                       # weights come from a checkpoint (hf_loader.load_llama_decoder)
            self.packed = self.codes = self.codebooks = self.scales = None
            return
        g = torch.Generator(device=dev)
        g.manual_seed(seed)
        # the FULL layer is generated and this rank keeps its rows: the model is the same for every world size
        full = torch.randint(-128, 128, (out_f, in_f // 8, 2), dtype=torch.int8, device=dev, generator=g)
        codes = full[self.begin:self.end].contiguous()
        self.packed = ops.code2x8_prepack(codes)
        self.codes = codes if keep_codes else None
        del full
        # scaled so activations stay O(1) through the stack: W entries ~ N(0, 2 * s^2), want var(y) ~ var(x)
        s = (1.0 / (2.0 * in_f)) ** 0.5
        self.codebooks = (torch.randn((2, 256, 1, 8), device=dev, generator=g) * s).to(torch.float16)
        self.scales = torch.ones(self.rows, dtype=torch.float16, device=dev)


class Decoder:
    def __init__(self, cfg: DecoderConfig, device, world: int = 1, rank: int = 0, group: Optional[ShardGroup] = None,
                 max_seq: int = 512, seed: int = 0, exchange: str = "push", pg=None, keep_codes: bool = False, grouped: bool = True, fuse_swiglu: bool = True,
                 skeleton: bool = False, dataflow: Optional[bool] = None):
        self.cfg, self.dev, self.world, self.rank, self.group, self.max_seq = cfg, torch.device(device), world, rank, group, max_seq
        self.exchange, self.pg, self.grouped = exchange, pg, grouped
        # LL exchange: GEMVs as dataflow launches (_step_ll_dataflow).  Off by default: measured slower
        self.dataflow = (os.environ.get("LUTVQ_LL_DATAFLOW", "0") != "0") if dataflow is None else bool(dataflow)
        self.fuse_swiglu = fuse_swiglu and grouped and not (world > 1 and exchange == "nccl")
        if (world > 1 and exchange == "push") or exchange == "ll":
            assert group is not None
        self.lib = _lib.lib()
        dev, H = self.dev, cfg.hidden
        shapes = cfg.linear_shapes()
        self.layers: List[dict] = []
        for li in range(cfg.n_layers):
            L = {name: _Linear(i, o, dev, seed * 100003 + li * 16 + j, world, rank, keep_codes, skeleton) for j, (name, (i, o)) in enumerate(shapes.items())}
            L["norm1"] = torch.ones(H, dtype=torch.float16, device=dev)
            L["norm2"] = torch.ones(H, dtype=torch.float16, device=dev)
            kv = cfg.n_kv_heads
            L["kcache"] = torch.zeros((kv, max_seq, cfg.head_dim), dtype=torch.float16, device=dev)
            L["vcache"] = torch.zeros((kv, max_seq, cfg.head_dim), dtype=torch.float16, device=dev)
            self.layers.append(L)
        g = torch.Generator(device=dev)
        g.manual_seed(seed + 7)
        self.embed = (torch.randn((cfg.vocab, H), device=dev, generator=g)).to(torch.float16)
        self.lm_head = (torch.randn((cfg.vocab, H), device=dev, generator=g) * H ** -0.5).to(torch.float16)
        self.final_norm = torch.ones(H, dtype=torch.float16, device=dev)
        # activations
        f16 = dict(dtype=torch.float16, device=dev)
        self.resid = torch.zeros(H, **f16)
        self.normed = torch.zeros(H, **f16)
        self.attn_out = torch.zeros(H, **f16)
        self.act = torch.zeros(cfg.intermediate, **f16)
        self.bufs = {name: torch.zeros(o, **f16) for name, (i, o) in shapes.items()}      # one-GPU outputs
        self.logits = torch.zeros(cfg.vocab, **f16)
        self.token = torch.zeros(1, dtype=torch.long, device=dev)
        self.pos = torch.zeros(1, dtype=torch.int32, device=dev)
        # workspace ring: every GEMV takes the next region (include/lutvq.h X_READY contract)
        max_rows = max(l.rows for l in self.layers[0].values() if isinstance(l, _Linear))
        self.ws_need = (3 * int(self.lib.lutvq_packed_workspace_bytes(max_rows, max(i for i, _ in shapes.values()))) + 255) // 256 * 256   # a group of <= 3
        self.ws = torch.zeros(ops.XREADY_WS_RING * self.ws_need, dtype=torch.uint8, device=dev)
        self._ws_i = 0
        if world > 1 and exchange == "nccl":
            self._nccl_local = {name: torch.zeros(-(-o // (32 * world)) * 32, **f16) for name, (i, o) in shapes.items()}
            self._nccl_all = {name: torch.zeros(world * self._nccl_local[name].numel(), **f16) for name in shapes}
        self.graph = None
        self.launches_per_token = 0

    # ---- building blocks ----
    def _st(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)

    def _next_ws(self):
        i = self._ws_i
        self._ws_i = (i + 1) % ops.XREADY_WS_RING
        return ctypes.c_void_p(self.ws.data_ptr() + i * self.ws_need)

    def _gemv(self, lin: _Linear, x: torch.Tensor, name: str, x_ready: bool, slot: Optional[int]):
        """launch; returns (output tensor, slot to wait on or None)"""
        P = lambda t: ctypes.c_void_p(t.data_ptr())
        flags = _lib.FLAG_X_READY if x_ready else 0
        self.launches_per_token += 1
        if self.world == 1:
            out = self.bufs[name]
            _lib.check(self.lib.lutvq_code2x8_matmat_packed(P(x), P(lin.packed), P(lin.codebooks), P(lin.scales), None, P(out), 1, lin.out_f,
                                                            lin.in_f, 0, flags, self._next_ws(), self.ws_need, self._st()))
            return out, None
        if self.exchange == "push":
            _lib.check(self.lib.lutvq_sharded_matmat_packed(self.group._h, slot, lin.begin, P(x), P(lin.packed), P(lin.codebooks), P(lin.scales),
                                                            None, lin.rows, lin.in_f, 0, flags, self._next_ws(), self.ws_need, self._st()))
            return self.group.slot_tensor(slot, lin.out_f, torch.float16), slot
        # NCCL all-gather baseline: local rows, then the collective
        import torch.distributed as dist
        loc, allb = self._nccl_local[name], self._nccl_all[name]
        _lib.check(self.lib.lutvq_code2x8_matmat_packed(P(x), P(lin.packed), P(lin.codebooks), P(lin.scales), None, P(loc), 1, lin.rows,
                                                        lin.in_f, 0, flags, self._next_ws(), self.ws_need, self._st()))
        dist.all_gather_into_tensor(allb, loc, group=self.pg)
        self.launches_per_token += 1
        per = loc.numel()
        if per * self.world == lin.out_f:
            return allb, None
        out = self.bufs[name]
        for r in range(self.world):
            b, e = shard_rows(lin.out_f, self.world, r)
            out[b:e].copy_(allb[r * per: r * per + (e - b)])
        return out, None

    def _wait(self, slot):
        if slot is not None:
            _lib.check(self.lib.lutvq_shard_wait(self.group._h, slot, self._st()))
            self.launches_per_token += 1

    def _group_gemv(self, L, names, x, gate=None):
        """the linears that share an input go out as ONE grouped launch (lutvq_*_matmat_group_packed) and, when sharded,
        ONE exchange launch.  None of this may use LUTVQ_FLAG_X_READY: under programmatic dependent launch the kernel
        after an X_READY kernel can start before the kernel that PRODUCED x (two launches back) has finished, so the
        promise only holds for inputs that were complete before the whole overlapped run began (include/lutvq.h)."""
        n = len(names)
        lins = [L[k] for k in names]
        if self.world > 1 and self.exchange == "nccl" or not self.grouped:
            outs, slots = [], []
            for k in names:
                slot = self.group.next_slot() if (self.world > 1 and self.exchange == "push") else None
                o, s = self._gemv(L[k], x, k, x_ready=False, slot=slot)
                outs.append(o)
                slots.append(s)
            for s in slots:
                self._wait(s)
            return outs
        arr = lambda ts: (ctypes.c_void_p * n)(*[None if t is None else t.data_ptr() for t in ts])
        P = lambda t: ctypes.c_void_p(t.data_ptr())
        self.launches_per_token += 1
        if self.world == 1:
            outs = [self.bufs[k] for k in names]
            pm = (ctypes.c_int * n)(*[l.out_f for l in lins])
            tail = (n, arr([l.packed for l in lins]), arr([l.codebooks for l in lins]), arr([l.scales for l in lins]), arr([None] * n),
                    arr(outs), pm, lins[0].in_f, 0, 0, self._next_ws(), self.ws_need, self._st())
            if gate is None:
                _lib.check(self.lib.lutvq_code2x8_matmat_group_packed(P(x), *tail))
            else:                                             # x = silu(gate) * up inside the kernel's activation load
                _lib.check(self.lib.lutvq_code2x8_matmat_group_packed_swiglu(P(gate), P(x), *tail))
            return outs
        slots = [self.group.next_slot() for _ in names]
        sl = (ctypes.c_int * n)(*slots)
        ro = (ctypes.c_size_t * n)(*[l.begin for l in lins])
        rl = (ctypes.c_int * n)(*[l.rows for l in lins])
        tail = (arr([l.packed for l in lins]), arr([l.codebooks for l in lins]), arr([l.scales for l in lins]), arr([None] * n), rl,
                lins[0].in_f, 0, 0, self._next_ws(), self.ws_need, self._st())
        if gate is None:
            _lib.check(self.lib.lutvq_sharded_matmat_group_packed(self.group._h, n, sl, ro, P(x), *tail))
        else:
            _lib.check(self.lib.lutvq_sharded_matmat_group_packed_swiglu(self.group._h, n, sl, ro, P(gate), P(x), *tail))
        mask = 0
        for s_ in slots:
            mask |= 1 << s_
        _lib.check(self.lib.lutvq_shard_wait_mask(self.group._h, mask, self._st()))
        self.launches_per_token += 1
        return [self.group.slot_tensor(s_, l.out_f, torch.float16) for s_, l in zip(slots, lins)]

    # ---- LL exchange mode: every sharded output goes out as LL words, every consumer polls them (no exchange launches) ----
    def _ll_next(self):
        s = self._ll_slot
        self._ll_slot = (s + 1) % self.group.n_slots
        self._ll_seq += 1
        return s, self._ll_seq

    def _group_ll(self, L, names, x=None, x_ll=None, gate_ll=None, dataflow=False):
        """-> [(slot, seq)] of the outputs; x: plain tensor, or x_ll / gate_ll = (slot, seq) of earlier exchanges.
        dataflow (x_ll only): the launch does not wait for the launch before it -- it polls its LL activations (lutvq.h)"""
        lins = [L[k] for k in names]
        outs = [self._ll_next() for _ in names]
        n = len(names)
        arr = lambda ts: (ctypes.c_void_p * n)(*[None if t is None else t.data_ptr() for t in ts])
        sl = (ctypes.c_int * n)(*[o[0] for o in outs])
        sq = (ctypes.c_uint * n)(*[o[1] for o in outs])
        ro = (ctypes.c_size_t * n)(*[l.begin for l in lins])
        rl = (ctypes.c_int * n)(*[l.rows for l in lins])
        xs, xq = x_ll if x_ll is not None else (-1, 0)
        gs, gq = gate_ll if gate_ll is not None else (-1, 0)
        _lib.check(self.lib.lutvq_sharded_matmat_group_ll(
            self.group._h, n, sl, ro, sq, None if x is None else ctypes.c_void_p(x.data_ptr()), xs, xq, gs, gq,
            arr([l.packed for l in lins]), arr([l.codebooks for l in lins]), arr([l.scales for l in lins]), arr([None] * n), rl,
            lins[0].in_f, 0, _lib.FLAG_X_READY if (dataflow and x_ll is not None) else 0, self._next_ws(), self.ws_need, self._st()))
        self.launches_per_token += 1
        return outs

    def _step_ll_dataflow(self):
        """_step_ll with every GEMV a dataflow launch: the glue kernels also write their outputs as LL words (local slots), so each
        GEMV streams its codes and builds its LUT as soon as the words arrive instead of after the kernel before it has completed
        (and the launch gap after that).  The glue kernels keep their dependency wait: completion stays transitive."""
        cfg, lib, P, st, h = self.cfg, self.lib, (lambda t: ctypes.c_void_p(t.data_ptr())), self._st, self.group._h
        self._ll_slot, self._ll_seq = 0, 0
        _lib.check(lib.lutvq_shard_ll_tick(h, st()))
        self.resid.copy_(torch.index_select(self.embed, 0, self.token)[0])
        delta = (-1, 0)
        for L in self.layers:
            n1 = self._ll_next()
            _lib.check(lib.lutvq_decode_add_rmsnorm_ll_out(h, P(self.resid), delta[0], delta[1], P(L["norm1"]), P(self.normed), n1[0], n1[1],
                                                           cfg.hidden, cfg.eps, st()))
            q, k, v = self._group_ll(L, ["q", "k", "v"], x_ll=n1, dataflow=True)
            ao = self._ll_next()
            _lib.check(lib.lutvq_decode_attention_ll_out(h, q[0], k[0], v[0], q[1], k[1], v[1], P(L["kcache"]), P(L["vcache"]), P(self.pos),
                                                         P(self.attn_out), ao[0], ao[1], cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, self.max_seq,
                                                         cfg.rope_theta, st()))
            (o,) = self._group_ll(L, ["o"], x_ll=ao, dataflow=True)
            n2 = self._ll_next()
            _lib.check(lib.lutvq_decode_add_rmsnorm_ll_out(h, P(self.resid), o[0], o[1], P(L["norm2"]), P(self.normed), n2[0], n2[1],
                                                           cfg.hidden, cfg.eps, st()))
            gate, up = self._group_ll(L, ["gate", "up"], x_ll=n2, dataflow=True)
            (delta,) = self._group_ll(L, ["down"], x_ll=up, gate_ll=gate, dataflow=True)
            self.launches_per_token += 3
        _lib.check(lib.lutvq_decode_add_rmsnorm_ll(h, P(self.resid), delta[0], delta[1], P(self.final_norm), P(self.normed), cfg.hidden, cfg.eps, st()))
        torch.mv(self.lm_head, self.normed, out=self.logits)
        self.token.copy_(torch.argmax(self.logits).reshape(1))
        self.pos.add_(1)
        self.launches_per_token += 3

    def _step_ll(self):
        cfg, lib, P, st, h = self.cfg, self.lib, (lambda t: ctypes.c_void_p(t.data_ptr())), self._st, self.group._h
        self._ll_slot, self._ll_seq = 0, 0
        _lib.check(lib.lutvq_shard_ll_tick(h, st()))
        self.resid.copy_(torch.index_select(self.embed, 0, self.token)[0])
        delta = None
        for L in self.layers:
            if delta is None:
                _lib.check(lib.lutvq_decode_add_rmsnorm(P(self.resid), None, P(L["norm1"]), P(self.normed), cfg.hidden, cfg.eps, st()))
            else:
                _lib.check(lib.lutvq_decode_add_rmsnorm_ll(h, P(self.resid), delta[0], delta[1], P(L["norm1"]), P(self.normed), cfg.hidden, cfg.eps, st()))
            q, k, v = self._group_ll(L, ["q", "k", "v"], x=self.normed)
            _lib.check(lib.lutvq_decode_attention_ll(h, q[0], k[0], v[0], q[1], k[1], v[1], P(L["kcache"]), P(L["vcache"]), P(self.pos),
                                                     P(self.attn_out), cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, self.max_seq, cfg.rope_theta, st()))
            (o,) = self._group_ll(L, ["o"], x=self.attn_out)
            _lib.check(lib.lutvq_decode_add_rmsnorm_ll(h, P(self.resid), o[0], o[1], P(L["norm2"]), P(self.normed), cfg.hidden, cfg.eps, st()))
            gate, up = self._group_ll(L, ["gate", "up"], x=self.normed)
            (delta,) = self._group_ll(L, ["down"], x_ll=up, gate_ll=gate)            # silu(gate) * up inside the activation load
            self.launches_per_token += 3
        _lib.check(lib.lutvq_decode_add_rmsnorm_ll(h, P(self.resid), delta[0], delta[1], P(self.final_norm), P(self.normed), cfg.hidden, cfg.eps, st()))
        torch.mv(self.lm_head, self.normed, out=self.logits)
        self.token.copy_(torch.argmax(self.logits).reshape(1))
        self.pos.add_(1)
        self.launches_per_token += 3

    def step(self):
        """one token: reads self.token / self.pos, writes self.logits and the next token"""
        if self.exchange == "ll":
            return self._step_ll_dataflow() if self.dataflow else self._step_ll()
        cfg, lib, P = self.cfg, self.lib, (lambda t: ctypes.c_void_p(t.data_ptr()))
        st = self._st
        self.resid.copy_(torch.index_select(self.embed, 0, self.token)[0])
        delta = None
        for L in self.layers:
            _lib.check(lib.lutvq_decode_add_rmsnorm(P(self.resid), None if delta is None else P(delta), P(L["norm1"]), P(self.normed), cfg.hidden, cfg.eps, st()))
            q, k, v = self._group_gemv(L, ["q", "k", "v"], self.normed)
            _lib.check(lib.lutvq_decode_attention(P(q), P(k), P(v), P(L["kcache"]), P(L["vcache"]), P(self.pos), P(self.attn_out), cfg.n_heads,
                                                  cfg.n_kv_heads, cfg.head_dim, self.max_seq, cfg.rope_theta, st()))
            (o,) = self._group_gemv(L, ["o"], self.attn_out)
            _lib.check(lib.lutvq_decode_add_rmsnorm(P(self.resid), P(o), P(L["norm2"]), P(self.normed), cfg.hidden, cfg.eps, st()))
            gate, up = self._group_gemv(L, ["gate", "up"], self.normed)
            if self.fuse_swiglu:                              # SwiGLU inside down_proj's activation load: one launch fewer
                (delta,) = self._group_gemv(L, ["down"], up, gate=gate)
                self.launches_per_token += 3
            else:
                _lib.check(lib.lutvq_decode_silu_mul(P(gate), P(up), P(self.act), cfg.intermediate, st()))
                (delta,) = self._group_gemv(L, ["down"], self.act)
                self.launches_per_token += 4
        _lib.check(lib.lutvq_decode_add_rmsnorm(P(self.resid), P(delta), P(self.final_norm), P(self.normed), cfg.hidden, cfg.eps, st()))
        torch.mv(self.lm_head, self.normed, out=self.logits)
        self.token.copy_(torch.argmax(self.logits).reshape(1))
        self.pos.add_(1)
        self.launches_per_token += 1

    def capture(self, stream: torch.cuda.Stream):
        """warm one eager step, then capture the step as a CUDA graph (one-GPU and push mode; NCCL mode stays eager)"""
        with torch.cuda.stream(stream):
            self.launches_per_token = 0
            self.step()
            n = self.launches_per_token
            torch.cuda.synchronize(self.dev)
            if self.world > 1 and self.exchange == "nccl":
                self.launches_per_token = n
                return
            self._barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                self.step()
            self.graph = g
            self.launches_per_token = n

    def _barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            torch.cuda.synchronize(self.dev)
            dist.barrier(group=self.pg)

    def reset(self, token: int = 1):
        self.token.fill_(token)
        self.pos.zero_()

    def run_token(self):
        if self.graph is not None:
            self.graph.replay()
        else:
            self.step()
