#!/usr/bin/env python
"""Golden vectors for the PQ4 fast-scan row -- TEST INFRASTRUCTURE.

Runs the REFERENCE'S OWN compiled code (oracle/_ref/libpq4_ref.so = benchmark/kernel/faiss-simd/pq4_fast_scan.cpp +
pq4_fast_scan_search_1.cpp + simdlib_avx2.h, built in place by `make -C oracle ref` against the header shims in
oracle/faiss_shim/) on seeded inputs and freezes inputs + outputs under tests/golden/pq4_*.npz.
Only runnable in the build container; the fixtures travel."""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
P, Z, I = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int

# (name, nq, bbs, M, ntotal, scale): every (nq, bbs/32) the reference instantiates appears once; odd M; ntotal not a multiple
# of bbs; nsq > 257 (uint16 wrap-around); a NormTableScaler that overflows
CASES = [("q1_b32_m32", 1, 32, 32, 100, -1), ("q1_b64_m16", 1, 64, 16, 200, -1), ("q1_b96_m8", 1, 96, 8, 96, -1),
         ("q1_b128_m64", 1, 128, 64, 300, -1), ("q1_b160_m600", 1, 160, 600, 320, -1), ("q2_b32_m7", 2, 32, 7, 64, -1),
         ("q2_b64_m16_s3", 2, 64, 16, 200, 3), ("q3_b32_m64_s200", 3, 32, 64, 96, 200), ("q4_b32_m20", 4, 32, 20, 33, -1)]


def _p(a):
    return a.ctypes.data_as(P)


def main(out_dir: str) -> None:
    import oracle as orc
    orc.build()
    ref = orc.ref_pq4()
    assert ref is not None, "oracle/_ref/libpq4_ref.so missing: make -C oracle ref"
    ref.ref_pq4_pack_codes.argtypes = [P, Z, Z, Z, Z, Z, P]
    ref.ref_pq4_pack_codes_range.argtypes = [P, Z, Z, Z, Z, Z, P]
    ref.ref_pq4_get_packed_element.argtypes = [P, Z, Z, Z, Z]
    ref.ref_pq4_set_packed_element.argtypes = [P, I, Z, Z, Z, Z]
    ref.ref_pq4_accumulate_loop.argtypes = [I, Z, I, I, P, P, P, I]
    ref.ref_pq4_code_packer.argtypes = [Z, Z, P, Z, P, P, P]
    os.makedirs(out_dir, exist_ok=True)
    for name, nq, bbs, M, ntotal, scale in CASES:
        rng = np.random.default_rng(sum(map(ord, name)))
        nsq, nb = orc.pq4_nsq(M), -(-ntotal // bbs) * bbs
        codes = rng.integers(0, 256, size=(ntotal, (M + 1) // 2), dtype=np.uint8)
        if M % 2:
            codes[:, -1] &= 15                     # an odd M leaves the last high nibble unused; faiss writes zeros there
        lut = rng.integers(0, 256, size=(nq, nsq, 16), dtype=np.uint8)
        blocks = orc.aligned_u8(nb * nsq // 2)
        assert ref.ref_pq4_pack_codes(_p(codes), ntotal, M, nb, bbs, nsq, _p(blocks)) == 0
        # the same image built by ranges (two pieces) must equal the one-shot packing
        cut = ntotal // 3 + 1
        blocks_r = orc.aligned_u8(nb * nsq // 2)
        assert ref.ref_pq4_pack_codes_range(_p(np.ascontiguousarray(codes[:cut])), M, 0, cut, bbs, nsq, _p(blocks_r)) == 0
        assert ref.ref_pq4_pack_codes_range(_p(np.ascontiguousarray(codes[cut:])), M, cut, ntotal, bbs, nsq, _p(blocks_r)) == 0
        assert np.array_equal(blocks, blocks_r)
        lut_packed = orc.aligned_u8(nq * nsq * 16)
        ref.ref_pq4_pack_LUT(nq, nsq, _p(lut), _p(lut_packed))
        accu = np.zeros((nq, nb), dtype=np.uint16)
        assert ref.ref_pq4_accumulate_loop(nq, nb, bbs, nsq, _p(blocks), _p(lut_packed), _p(accu), scale) == 0
        # element getter over a sample of (vector, sq)
        vs, sqs = rng.integers(0, ntotal, 64), rng.integers(0, nsq, 64)
        elems = np.array([ref.ref_pq4_get_packed_element(_p(blocks), bbs, nsq, int(v), int(s)) for v, s in zip(vs, sqs)], dtype=np.uint8)
        np.savez_compressed(os.path.join(out_dir, f"pq4_{name}.npz"), cfg=np.array([nq, bbs, M, ntotal, scale], dtype=np.int64), codes=codes,
                            lut=lut, blocks=np.array(blocks), lut_packed=np.array(lut_packed), accu=accu, elem_v=vs, elem_sq=sqs, elems=elems)
        print(name, "accu[0,:4] =", accu[0, :4])
    # qbs LUT packing (with and without a query map), the code packer and set_packed_element
    rng = np.random.default_rng(77)
    nsq = 8
    lut = rng.integers(0, 256, size=(9, nsq, 16), dtype=np.uint8)
    qbs = 0x1323                                                     # blocks of 3, 2, 3, 1 queries
    d1 = orc.aligned_u8(9 * nsq * 16)
    assert ref.ref_pq4_pack_LUT_qbs(qbs, nsq, _p(lut), _p(d1)) == 9
    q_map = rng.permutation(9).astype(np.int32)
    d2 = orc.aligned_u8(9 * nsq * 16)
    assert ref.ref_pq4_pack_LUT_qbs_q_map(qbs, nsq, _p(lut), _p(q_map), _p(d2)) == 9
    bbs = 64
    flat = rng.integers(0, 256, size=(5, nsq // 2), dtype=np.uint8)
    offsets = np.array([0, 17, 63, 64 + 9, 3 * 64 + 40], dtype=np.int64)
    block = orc.aligned_u8(4 * (nsq // 2) * bbs)
    back = np.zeros_like(flat)
    sizes = np.zeros(3, dtype=np.uint64)
    for k in range(5):
        ref.ref_pq4_code_packer(nsq, bbs, _p(np.ascontiguousarray(flat[k])), int(offsets[k]), _p(block), _p(back[k]), _p(sizes))
    assert np.array_equal(back, flat)
    np.savez_compressed(os.path.join(out_dir, "pq4_packers.npz"), lut=lut, qbs=np.array([qbs]), lut_qbs=np.array(d1), q_map=q_map,
                        lut_qbs_qmap=np.array(d2), flat=flat, offsets=offsets, block=np.array(block), sizes=sizes, nsq=np.array([nsq]), bbs=np.array([bbs]))
    print("packers ok", sizes)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "..", "tests", "golden"))
