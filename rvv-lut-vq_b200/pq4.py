"""PQ4 fast-scan on the GPU: host-side mirror of ``benchmark/kernel/faiss-simd/pq4_fast_scan.h``.

Same function names, argument meaning and packed layouts as the reference's (faiss-derived) API; every array is a CUDA
uint8 / uint16 tensor and the work is done by csrc/pq4.cu through the C-ABI (no CPU path).  Differences, all stated in
include/lutvq.h: results go to a uint16 matrix (the reference's store-to-memory result handler) instead of a callback object;
``scaler`` is the integer of NormTableScaler (None = no scaler); (nq, bbs) may be any nq in 1..4 with any bbs % 32 == 0."""
from __future__ import annotations

from typing import Optional

import torch

from ._lib import check, lib
from .ops import _ptr, _require_cuda, _stream


def _u8(t: torch.Tensor, name: str) -> torch.Tensor:
    assert t.dtype == torch.uint8 and t.is_contiguous(), f"{name} must be a contiguous uint8 tensor"
    return t


def aligned_empty(n: int, device, dtype=torch.uint8) -> torch.Tensor:
    """zeroed tensor whose data pointer is 32-byte aligned (the accumulation checks it like the reference does)"""
    t = torch.zeros(n, dtype=dtype, device=device)          # the CUDA caching allocator hands out 512-byte aligned blocks
    assert t.data_ptr() % 32 == 0
    return t


def pq4_nsq(M: int) -> int:
    """number of sub-quantizers rounded up to a multiple of 2 (pq4_fast_scan.h:37)"""
    return (M + 1) // 2 * 2


def pq4_pack_codes(codes: torch.Tensor, ntotal: int, M: int, nb: int, bbs: int, nsq: int, blocks: Optional[torch.Tensor] = None) -> torch.Tensor:
    """pq4_fast_scan.cpp:45-91.  codes [ntotal, (M+1)//2] uint8 -> blocks uint8[nb * nsq // 2] (unused bytes 0)."""
    dev = _require_cuda(codes)
    _u8(codes, "codes")
    assert codes.numel() >= ntotal * ((M + 1) // 2)
    if blocks is None:
        blocks = aligned_empty(nb * nsq // 2, dev)
    assert _u8(blocks, "blocks").numel() >= nb * nsq // 2
    with torch.cuda.device(dev):
        check(lib().lutvq_pq4_pack_codes(_ptr(codes), ntotal, M, nb, bbs, nsq, _ptr(blocks), _stream(dev)))
    return blocks


def pq4_pack_codes_range(codes: torch.Tensor, M: int, i0: int, i1: int, bbs: int, nsq: int, blocks: torch.Tensor) -> torch.Tensor:
    """pq4_fast_scan.cpp:93-137: ORs vectors [i0, i1) (codes [i1 - i0, (M+1)//2]) into `blocks`, the rest untouched."""
    dev = _require_cuda(codes, blocks)
    assert _u8(codes, "codes").numel() >= (i1 - i0) * ((M + 1) // 2)
    assert _u8(blocks, "blocks").numel() >= -(-i1 // bbs) * bbs * nsq // 2
    with torch.cuda.device(dev):
        check(lib().lutvq_pq4_pack_codes_range(_ptr(codes), M, i0, i1, bbs, nsq, _ptr(blocks), _stream(dev)))
    return blocks


def _vector_address(bbs: int, vector_id: int, sq: int):
    """pq4_fast_scan.cpp:144-166 get_vector_specific_address -> (byte offset inside the block, high nibble?)"""
    vector_id %= bbs
    shift = vector_id > 15
    vector_id &= 15
    address = vector_id << 1 if vector_id < 8 else ((vector_id - 8) << 1) + 1
    if sq & 1:
        address += 16
    return (sq >> 1) * bbs + address, shift


def pq4_get_packed_element(data: torch.Tensor, bbs: int, nsq: int, vector_id: int, sq: int) -> int:
    """pq4_fast_scan.cpp:168-184 (one device read: a debugging accessor, as in the reference)"""
    base = (vector_id // bbs) * (((nsq + 1) // 2) * bbs)
    address, shift = _vector_address(bbs, vector_id, sq)
    v = int(data[base + address].item())
    return v >> 4 if shift else v & 15


def pq4_set_packed_element(data: torch.Tensor, code: int, bbs: int, nsq: int, vector_id: int, sq: int) -> None:
    """pq4_fast_scan.cpp:186-204"""
    base = (vector_id // bbs) * (((nsq + 1) // 2) * bbs)
    address, shift = _vector_address(bbs, vector_id, sq)
    v = int(data[base + address].item())
    data[base + address] = ((code << 4) | (v & 15)) & 255 if shift else (code | (v & ~15)) & 255


class CodePackerPQ4:
    """pq4_fast_scan.cpp:210-248: packs / unpacks ONE flat code (nsq / 2 bytes, two sub-quantizers per byte) at a time."""

    def __init__(self, nsq: int, bbs: int):
        self.nsq, self.nvec = nsq, bbs
        self.code_size = (nsq * 4 + 7) // 8
        self.block_size = ((nsq + 1) // 2) * bbs

    def pack_1(self, flat_code, offset: int, block: torch.Tensor) -> None:
        for i in range(self.code_size):
            code = int(flat_code[i])
            pq4_set_packed_element(block, code & 15, self.nvec, self.nsq, offset, 2 * i)
            pq4_set_packed_element(block, code >> 4, self.nvec, self.nsq, offset, 2 * i + 1)

    def unpack_1(self, block: torch.Tensor, offset: int):
        return [pq4_get_packed_element(block, self.nvec, self.nsq, offset, 2 * i) | (pq4_get_packed_element(block, self.nvec, self.nsq, offset, 2 * i + 1) << 4)
                for i in range(self.code_size)]


def pq4_pack_LUT(nq: int, nsq: int, src: torch.Tensor, dest: Optional[torch.Tensor] = None) -> torch.Tensor:
    """pq4_fast_scan.cpp:253-264.  src [nq, nsq, 16] uint8 -> dest uint8[nq * nsq * 16]"""
    dev = _require_cuda(src)
    assert _u8(src, "src").numel() >= nq * nsq * 16
    if dest is None:
        dest = aligned_empty(nq * nsq * 16, dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_pq4_pack_LUT(nq, nsq, _ptr(src), _ptr(_u8(dest, "dest")), _stream(dev)))
    return dest


def pq4_qbs_to_nq(qbs: int) -> int:
    """sum of the base-16 digits (pq4_fast_scan.h:146)"""
    return lib().lutvq_pq4_qbs_to_nq(qbs)


def pq4_pack_LUT_qbs(qbs: int, nsq: int, src: torch.Tensor, dest: Optional[torch.Tensor] = None) -> torch.Tensor:
    """pq4_fast_scan.cpp:266-281: one pq4_pack_LUT per digit of qbs"""
    return pq4_pack_LUT_qbs_q_map(qbs, nsq, src, None, dest)


def pq4_pack_LUT_qbs_q_map(qbs: int, nsq: int, src: torch.Tensor, q_map: Optional[torch.Tensor], dest: Optional[torch.Tensor] = None) -> torch.Tensor:
    """pq4_fast_scan.cpp:305-322: like pq4_pack_LUT_qbs with the source rows remapped through q_map (int32, on the device)"""
    dev = _require_cuda(src, q_map)
    nq = pq4_qbs_to_nq(qbs)
    if q_map is not None:
        assert q_map.dtype == torch.int32 and q_map.is_contiguous() and q_map.numel() >= nq
    else:
        assert _u8(src, "src").numel() >= nq * nsq * 16
    if dest is None:
        dest = aligned_empty(nq * nsq * 16, dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_pq4_pack_LUT_qbs(qbs, nsq, _ptr(_u8(src, "src")), _ptr(q_map), _ptr(_u8(dest, "dest")), _stream(dev)))
    return dest


def _accu(accu: Optional[torch.Tensor], nq: int, nb: int, dev) -> torch.Tensor:
    if accu is None:
        return torch.empty((nq, nb), dtype=torch.uint16, device=dev)
    assert accu.dtype == torch.uint16 and accu.dim() == 2 and accu.shape[0] >= nq and accu.shape[1] >= nb and accu.stride(1) == 1
    return accu


def pq4_accumulate_loop(nq: int, nb: int, bbs: int, nsq: int, codes: torch.Tensor, LUT: torch.Tensor, accu: Optional[torch.Tensor] = None,
                        scaler: Optional[int] = None) -> torch.Tensor:
    """pq4_fast_scan_search_1.cpp:137-222 with a store-to-memory handler: -> uint16 [nq, nb] distances (sums modulo 2**16)."""
    dev = _require_cuda(codes, LUT)
    assert _u8(codes, "codes").numel() >= nb * nsq // 2 and _u8(LUT, "LUT").numel() >= nq * nsq * 16
    accu = _accu(accu, nq, nb, dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_pq4_accumulate_loop(nq, nb, bbs, nsq, _ptr(codes), _ptr(LUT), _ptr(accu), accu.stride(0), -1 if scaler is None else int(scaler),
                                              _stream(dev)))
    return accu


def pq4_accumulate_loop_qbs(qbs: int, nb: int, nsq: int, codes: torch.Tensor, LUT: torch.Tensor, accu: Optional[torch.Tensor] = None,
                            scaler: Optional[int] = None) -> torch.Tensor:
    """pq4_fast_scan.h:170-188: bbs = 32, query blocks given by the digits of qbs, LUT packed by pq4_pack_LUT_qbs."""
    dev = _require_cuda(codes, LUT)
    nq = pq4_qbs_to_nq(qbs)
    assert _u8(codes, "codes").numel() >= nb * nsq // 2 and _u8(LUT, "LUT").numel() >= nq * nsq * 16
    accu = _accu(accu, nq, nb, dev)
    with torch.cuda.device(dev):
        check(lib().lutvq_pq4_accumulate_loop_qbs(qbs, nb, nsq, _ptr(codes), _ptr(LUT), _ptr(accu), accu.stride(0), -1 if scaler is None else int(scaler),
                                                  _stream(dev)))
    return accu


def accumulate_to_mem(nq: int, ntotal2: int, nsq: int, codes: torch.Tensor, LUT: torch.Tensor, accu: Optional[torch.Tensor] = None) -> torch.Tensor:
    """pq4_fast_scan.h:190-207: nq (1..4) queries against ntotal2 (multiple of 32) vectors, no scaler, results to memory."""
    return pq4_accumulate_loop(nq, ntotal2, 32, nsq, codes, LUT, accu, None)
