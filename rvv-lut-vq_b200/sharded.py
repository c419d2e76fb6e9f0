"""Output-channel-sharded LUT-GEMV across the GPUs of one node (BASELINE.json configs[4]).

One process per GPU (torch.distributed is the plumbing).  Rank r of P owns rows ``shard_rows(O, P, r)`` of every
layer's codes / scales / bias; x and the 8 KiB codebooks are replicated.  Two exchange modes for the step that turns
the P output slices back into the whole y every rank needs as the next input:

  ``push``  (product) the GEMV fills this rank's slice of a symmetric gather slot; ONE exchange kernel
            (``ShardGroup.wait`` / ``wait_mask``) forwards the slice to every peer with coalesced NVLink stores, publishes
            the slot's epoch and spins (bounded) on local memory until all P slices have landed -- no NCCL, graph-capturable.
            ``fused_epilogue=True`` pushes from the GEMV's own epilogue instead (measured slower: DESIGN.md 4.5).
  ``nccl``  (baseline) local GEMV, then ``dist.all_gather_into_tensor`` of the slices.

The pure-host parts (row partition, handle exchange, slot rotation) are exercised on CPU with gloo in
tests/test_sharded_host.py; the device protocol on one GPU (two ranks in one process) and on 2+ GPUs in tests/test_gpu_sharded.py.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Tuple

import numpy as np

HANDLE_BYTES = 64           # include/lutvq.h LUTVQ_SHARD_HANDLE_BYTES


def shard_rows(out_features: int, world: int, rank: int, align: int = 32) -> Tuple[int, int]:
    """[begin, end) of the output rows rank owns: contiguous, whole 32-row sets (the GEMV's row-set granule) except
    possibly the last rank's tail; every row owned exactly once."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} not in [0, {world})")
    sets = -(-out_features // align)
    per, rem = divmod(sets, world)
    b = rank * per + min(rank, rem)
    e = b + per + (1 if rank < rem else 0)
    return min(b * align, out_features), min(e * align, out_features)


def exchange_handles(local_handle: bytes, group=None) -> bytes:
    """all-gather of the 64-byte IPC handles through torch.distributed (any backend: a CPU uint8 tensor on gloo, a CUDA
    one on nccl); returns the concatenation indexed by rank."""
    import torch
    import torch.distributed as dist
    assert len(local_handle) == HANDLE_BYTES
    world = dist.get_world_size(group)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    mine = torch.frombuffer(bytearray(local_handle), dtype=torch.uint8).to(dev)
    out = torch.empty(world * HANDLE_BYTES, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(out, mine, group=group)
    return bytes(out.cpu().numpy().tobytes())


class _CudaView:
    """zero-copy torch view of device memory owned by liblutvq (via __cuda_array_interface__)"""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3}
        self._owner = owner


class ShardGroup:
    """The symmetric gather buffer of one rank + its mapping of the peers' buffers."""

    def __init__(self, world: int, rank: int, max_out_features: int, n_slots: int = 4, device=None):
        import torch
        from . import _lib
        self._lib, self._check = _lib.lib(), _lib.check
        self.world, self.rank, self.n_slots = world, rank, n_slots
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self._h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            self._check(self._lib.lutvq_shard_create(ctypes.byref(self._h), world, rank, n_slots, max_out_features * 2))
        self._next_slot = 0
        self._views = {}

    # ---- setup ----
    def export_handle(self) -> bytes:
        buf = (ctypes.c_ubyte * HANDLE_BYTES)()
        self._check(self._lib.lutvq_shard_export(self._h, buf))
        return bytes(buf)

    def attach(self, handles: bytes) -> None:
        assert len(handles) == self.world * HANDLE_BYTES
        buf = (ctypes.c_ubyte * len(handles)).from_buffer_copy(handles)
        self._check(self._lib.lutvq_shard_attach(self._h, buf))

    def attach_same_process(self, groups: List["ShardGroup"]) -> None:
        arr = (ctypes.c_void_p * self.world)(*[g.base_ptr() for g in groups])
        self._check(self._lib.lutvq_shard_attach_ptrs(self._h, arr))

    @classmethod
    def from_process_group(cls, max_out_features: int, n_slots: int = 4, group=None) -> "ShardGroup":
        import torch.distributed as dist
        g = cls(dist.get_world_size(group), dist.get_rank(group), max_out_features, n_slots)
        if g.world > 1:
            g.attach(exchange_handles(g.export_handle(), group))
            dist.barrier(group)
        return g

    def base_ptr(self) -> int:
        return int(self._lib.lutvq_shard_base(self._h))

    # ---- per call ----
    def next_slot(self) -> int:
        s = self._next_slot
        self._next_slot = (s + 1) % self.n_slots
        return s

    def slot_tensor(self, slot: int, out_features: int, dtype):
        import torch
        key = (slot, out_features, dtype)
        v = self._views.get(key)
        if v is None:
            ptr = int(self._lib.lutvq_shard_slot(self._h, slot))
            raw = torch.as_tensor(_CudaView(ptr, out_features * 2, self), device=self.device)
            v = raw.view(dtype)
            self._views[key] = v
        return v

    def matmat(self, slot: int, row_offset: int, x, packed, codebooks, scales, bias, rows_local: int, *, x_ready: bool = False,
               fused_epilogue: bool = False) -> None:
        import torch
        from . import ops, _lib
        dtype = ops.check_use_bfloat16(x)
        in_features = x.size(-1)
        assert x.numel() == in_features, "sharded decode is batch 1"
        flags = (_lib.FLAG_X_READY if x_ready else 0) | (_lib.FLAG_PUSH_EPILOGUE if fused_epilogue else 0)
        with torch.cuda.device(self.device):
            ws = ops.workspace(self.device, rows_local, in_features, rotate=x_ready)
            self._check(self._lib.lutvq_sharded_matmat_packed(
                self._h, slot, row_offset, ops._ptr(x.contiguous()), ops._ptr(packed), ops._ptr(codebooks), ops._ptr(scales),
                ops._ptr(bias), rows_local, in_features, dtype, flags, ops._ptr(ws), ws.numel(), ops._stream(self.device)))

    def wait(self, slot: int) -> None:
        import torch
        from . import ops
        with torch.cuda.device(self.device):
            self._check(self._lib.lutvq_shard_wait(self._h, slot, ops._stream(self.device)))

    def wait_mask(self, mask: int) -> None:
        import torch
        from . import ops
        with torch.cuda.device(self.device):
            self._check(self._lib.lutvq_shard_wait_mask(self._h, mask, ops._stream(self.device)))

    # ---- LL exchange (include/lutvq.h '"LL" exchange'): no exchange launch, no flags; consumers poll the words they use ----
    def ll_tick(self) -> None:
        """++step counter; once per step (token) on every rank, before the step's first exchange"""
        from . import ops
        self._check(self._lib.lutvq_shard_ll_tick(self._h, ops._stream(self.device)))

    def matmat_group_ll(self, out_slots, row_offsets, out_seq, packed, codebooks, scales, bias, rows_local, in_features: int, *,
                        x=None, x_slot: int = -1, x_seq: int = 0, gate_slot: int = -1, gate_seq: int = 0, dtype=None, ws=None) -> None:
        """n <= 3 linears sharing an input in one launch; outputs as LL words into every rank's slot out_slots[i]"""
        import torch
        from . import ops
        n = len(packed)
        dt = ops.check_use_bfloat16(x) if x is not None else ops._DT[dtype or torch.float16]
        arr = lambda ts: (ctypes.c_void_p * n)(*[None if t is None else t.data_ptr() for t in ts])
        sl = (ctypes.c_int * n)(*out_slots)
        ro = (ctypes.c_size_t * n)(*row_offsets)
        sq = (ctypes.c_uint * n)(*out_seq)
        rl = (ctypes.c_int * n)(*rows_local)
        if ws is None:
            need = int(self._lib.lutvq_group_workspace_bytes(n, rl, in_features))
            key = ("ll", self.device.index)
            ws = self._views.get(key)
            if ws is None or ws.numel() < need:
                ws = torch.zeros(max(need, 1 << 18), dtype=torch.uint8, device=self.device)
                self._views[key] = ws
        with torch.cuda.device(self.device):
            self._check(self._lib.lutvq_sharded_matmat_group_ll(
                self._h, n, sl, ro, sq, None if x is None else ops._ptr(x.contiguous()), x_slot, x_seq, gate_slot, gate_seq,
                arr(packed), arr(codebooks), arr(scales), arr(bias if bias is not None else [None] * n), rl, in_features, dt, 0,
                ops._ptr(ws), ws.numel(), ops._stream(self.device)))

    def ll_unpack(self, slot: int, seq: int, out_features: int, dtype=None):
        """LL slot -> plain vector (waits, bounded, until every word of the exchange `seq` of this step has landed)"""
        import torch
        from . import ops
        y = torch.empty(out_features, dtype=dtype or torch.float16, device=self.device)
        with torch.cuda.device(self.device):
            self._check(self._lib.lutvq_shard_ll_unpack(self._h, slot, seq, ops._ptr(y), out_features, ops._stream(self.device)))
        return y

    def error(self) -> int:
        e = ctypes.c_uint(0)
        self._check(self._lib.lutvq_shard_error(self._h, ctypes.byref(e)))
        return int(e.value)

    def close(self) -> None:
        if self._h:
            self._views.clear()
            self._lib.lutvq_shard_destroy(self._h)
            self._h = ctypes.c_void_p()


class ShardedQuantizedLinear:
    """One rank's shard of an AQLM 2x8 linear layer.  ``forward(x)`` returns the WHOLE output vector [out_features]
    (a view of the gather slot in push mode), identical on every rank."""

    def __init__(self, full_or_local_codes, codebooks, scales, bias, out_features: int, world: int, rank: int,
                 group: Optional[ShardGroup] = None, mode: str = "push", codes_are_local: bool = False, pg=None):
        import torch
        from . import ops
        self.out_features, self.world, self.rank, self.mode, self.pg = out_features, world, rank, mode, pg
        self.begin, self.end = shard_rows(out_features, world, rank)
        self.rows = self.end - self.begin
        codes = full_or_local_codes if codes_are_local else full_or_local_codes[self.begin:self.end]
        assert codes.size(0) == self.rows
        self.in_features = codes.size(1) * 8
        self.codebooks = codebooks.contiguous()
        sl = (lambda t: None if t is None else t.reshape(-1)[(slice(None) if codes_are_local else slice(self.begin, self.end))].contiguous())
        self.scales, self.bias = sl(scales), sl(bias)
        self.packed = ops.code2x8_prepack(codes.contiguous())
        self.group = group
        if mode == "push":
            assert group is not None, "push mode needs a ShardGroup"
        elif mode == "nccl":
            # all_gather_into_tensor needs equal slices: pad every rank's slice to the largest
            self.max_rows = max(shard_rows(out_features, world, r)[1] - shard_rows(out_features, world, r)[0] for r in range(world))
            self._gathered = torch.empty(world * self.max_rows, dtype=codebooks.dtype, device=codes.device)
            self._local = torch.zeros(self.max_rows, dtype=codebooks.dtype, device=codes.device)
        else:
            raise ValueError(mode)

    def forward(self, x, *, x_ready: bool = False, slot: Optional[int] = None):
        import torch
        from . import ops
        if self.mode == "push":
            s = self.group.next_slot() if slot is None else slot
            self.group.matmat(s, self.begin, x, self.packed, self.codebooks, self.scales, self.bias, self.rows, x_ready=x_ready)
            self.group.wait(s)
            return self.group.slot_tensor(s, self.out_features, x.dtype)
        import torch.distributed as dist
        y = ops.code2x8_matmat_packed(x.reshape(1, -1), self.packed, self.rows, self.codebooks, self.scales, self.bias, x_ready=x_ready)
        self._local[: self.rows] = y.reshape(-1)
        if self.world == 1:
            return self._local[: self.rows]
        dist.all_gather_into_tensor(self._gathered, self._local, group=self.pg)
        if self.max_rows * self.world == self.out_features:
            return self._gathered
        parts = []
        for r in range(self.world):
            b, e = shard_rows(self.out_features, self.world, r)
            parts.append(self._gathered[r * self.max_rows: r * self.max_rows + (e - b)])
        return torch.cat(parts)

    __call__ = forward
