"""GPU parity of the PQ4 fast-scan entry points (csrc/pq4.cu) through the C-ABI: packed images and uint16
distances BIT-EXACT against the fixtures the reference's own compiled code produced (oracle/gen_golden_pq4.py), against the
oracle on seeded inputs, and -- at database sizes the oracle does not finish quickly -- through size-independent properties
(linearity of the sums in the table, invariance under the block size, pack -> get_packed_element round trip)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
PQ4_CASES = ["q1_b32_m32", "q1_b64_m16", "q1_b96_m8", "q1_b128_m64", "q1_b160_m600", "q2_b32_m7", "q2_b64_m16_s3", "q3_b32_m64_s200", "q4_b32_m20"]


@pytest.fixture(scope="module")
def pq4():
    import rvv_lut_vq_b200 as r
    assert os.path.exists(r._lib.LIB_PATH), "liblutvq.so must be prebuilt in-tree (no fallback)"
    assert torch.cuda.is_available()
    from rvv_lut_vq_b200 import pq4
    return pq4


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def host(t):
    return t.cpu().numpy()


@pytest.mark.parametrize("name", PQ4_CASES)
def test_pq4_fixture_bit_exact(pq4, golden_dir, name):
    g = dict(np.load(os.path.join(golden_dir, "pq4_" + name + ".npz")))
    nq, bbs, M, ntotal, scale = [int(v) for v in g["cfg"]]
    nsq, nb = pq4.pq4_nsq(M), -(-ntotal // bbs) * bbs
    codes = cu(g["codes"])
    blocks = pq4.pq4_pack_codes(codes, ntotal, M, nb, bbs, nsq)
    assert np.array_equal(host(blocks), g["blocks"])
    cut = ntotal // 3 + 1
    br = pq4.aligned_empty(nb * nsq // 2, "cuda")
    pq4.pq4_pack_codes_range(codes[:cut].contiguous(), M, 0, cut, bbs, nsq, br)
    pq4.pq4_pack_codes_range(codes[cut:].contiguous(), M, cut, ntotal, bbs, nsq, br)
    assert np.array_equal(host(br), g["blocks"])
    lp = pq4.pq4_pack_LUT(nq, nsq, cu(g["lut"]))
    assert np.array_equal(host(lp), g["lut_packed"])
    from rvv_lut_vq_b200._lib import lib
    for mode in (0, 1, 2):                       # size-based choice, PRMT kernel, byte-table kernel (where it applies)
        lib().lutvq_debug_set_pq4_mode(mode)
        try:
            accu = pq4.pq4_accumulate_loop(nq, nb, bbs, nsq, blocks, lp, None, None if scale < 0 else scale)
        finally:
            lib().lutvq_debug_set_pq4_mode(0)
        assert np.array_equal(host(accu), g["accu"]), f"mode {mode}"
    got = [pq4.pq4_get_packed_element(blocks, bbs, nsq, int(v), int(s)) for v, s in zip(g["elem_v"][:8], g["elem_sq"][:8])]
    assert got == g["elems"][:8].tolist()


def test_pq4_packers_fixture(pq4, golden_dir):
    g = dict(np.load(os.path.join(golden_dir, "pq4_packers.npz")))
    nsq, bbs, qbs = int(g["nsq"][0]), int(g["bbs"][0]), int(g["qbs"][0])
    assert pq4.pq4_qbs_to_nq(qbs) == 9
    assert np.array_equal(host(pq4.pq4_pack_LUT_qbs(qbs, nsq, cu(g["lut"]))), g["lut_qbs"])
    assert np.array_equal(host(pq4.pq4_pack_LUT_qbs_q_map(qbs, nsq, cu(g["lut"]), cu(g["q_map"]))), g["lut_qbs_qmap"])
    packer = pq4.CodePackerPQ4(nsq, bbs)
    assert [packer.code_size, packer.nvec, packer.block_size] == g["sizes"].tolist()
    block = pq4.aligned_empty(g["block"].size, "cuda")
    for flat, off in zip(g["flat"], g["offsets"]):
        packer.pack_1(flat, int(off), block)
    assert np.array_equal(host(block), g["block"])
    for flat, off in zip(g["flat"], g["offsets"]):
        assert packer.unpack_1(block, int(off)) == flat.tolist()


@pytest.mark.parametrize("nq,bbs,M,ntotal,scale", [(1, 32, 32, 4096, -1), (2, 256, 16, 5000, -1), (3, 64, 64, 3000, 7), (4, 32, 128, 2048, -1),
                                                  (4, 512, 8, 10000, 65535), (1, 32, 2, 32, 5), (2, 32, 1030, 96, -1)])
def test_pq4_vs_oracle(pq4, orc, nq, bbs, M, ntotal, scale):
    """seeded inputs, every nq with block sizes the reference's table does not instantiate; nsq > 514 crosses two segments"""
    rng = np.random.default_rng(nq * 1000 + bbs + M)
    nsq, nb = pq4.pq4_nsq(M), -(-ntotal // bbs) * bbs
    codes = rng.integers(0, 256, size=(ntotal, (M + 1) // 2), dtype=np.uint8)
    lut = rng.integers(0, 256, size=(nq, nsq, 16), dtype=np.uint8)
    blocks = pq4.pq4_pack_codes(cu(codes), ntotal, M, nb, bbs, nsq)
    assert np.array_equal(host(blocks), orc.c_pq4_pack_codes(codes, M, nb, bbs))
    lp = pq4.pq4_pack_LUT(nq, nsq, cu(lut))
    assert np.array_equal(host(lp), orc.c_pq4_pack_LUT(lut))
    accu = host(pq4.pq4_accumulate_loop(nq, nb, bbs, nsq, blocks, lp, None, None if scale < 0 else scale))
    want = orc.np_pq4_distances(codes, lut, M, scale)
    assert np.array_equal(accu[:, :ntotal], want)
    if nb > ntotal:                                              # padding vectors carry code 0 everywhere
        pad = orc.np_pq4_distances(np.zeros((1, (M + 1) // 2), np.uint8), lut, M, scale)
        assert np.array_equal(accu[:, ntotal:], np.repeat(pad, nb - ntotal, axis=1))


@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("nq,bbs,M,ntotal,scale", [(1, 32, 32, 2048, -1), (2, 64, 40, 1000, -1), (4, 32, 64, 4096, 31), (1, 128, 512, 640, -1),
                                                  (2, 32, 96, 512, 65535), (3, 96, 34, 960, 3), (1, 32, 300, 256, 1000)])
def test_pq4_both_kernels_vs_oracle(pq4, orc, mode, nq, bbs, M, ntotal, scale):
    """the PRMT kernel (mode 1) and the byte-table kernel (mode 2: >= 16 pairs, tables in shared memory) forced in turn:
    table counts that are not multiples of 32 (rotated walk wraps), 3 shared-memory planes, nsq * 255 >= 2^16 (per-half adds),
    scalers that overflow uint16"""
    from rvv_lut_vq_b200._lib import lib
    rng = np.random.default_rng(nq * 1000 + bbs + M + 7)
    nsq, nb = pq4.pq4_nsq(M), -(-ntotal // bbs) * bbs
    codes = rng.integers(0, 256, size=(ntotal, (M + 1) // 2), dtype=np.uint8)
    lut = rng.integers(0, 256, size=(nq, nsq, 16), dtype=np.uint8)
    blocks = pq4.pq4_pack_codes(cu(codes), ntotal, M, nb, bbs, nsq)
    lp = pq4.pq4_pack_LUT(nq, nsq, cu(lut))
    lib().lutvq_debug_set_pq4_mode(mode)
    try:
        accu = host(pq4.pq4_accumulate_loop(nq, nb, bbs, nsq, blocks, lp, None, None if scale < 0 else scale))
    finally:
        lib().lutvq_debug_set_pq4_mode(0)
    assert np.array_equal(accu[:, :ntotal], orc.np_pq4_distances(codes, lut, M, scale))


def test_pq4_qbs_blocks(pq4, orc):
    """pq4_accumulate_loop_qbs: blocks of 3, 2, 3, 1 queries, rows in block order"""
    rng = np.random.default_rng(5)
    M, ntotal, qbs = 24, 1000, 0x1323
    nsq, nb = 24, 1024
    codes = rng.integers(0, 256, size=(ntotal, M // 2), dtype=np.uint8)
    lut = rng.integers(0, 256, size=(9, nsq, 16), dtype=np.uint8)
    blocks = pq4.pq4_pack_codes(cu(codes), ntotal, M, nb, 32, nsq)
    lp = pq4.pq4_pack_LUT_qbs(qbs, nsq, cu(lut))
    assert np.array_equal(host(lp), orc.c_pq4_pack_LUT_qbs(qbs, lut))
    for scale in (None, 9):
        accu = host(pq4.pq4_accumulate_loop_qbs(qbs, nb, nsq, blocks, lp, None, scale))
        assert np.array_equal(accu[:, :ntotal], orc.np_pq4_distances(codes, lut, M, -1 if scale is None else scale))


def test_pq4_full_size_properties(pq4, orc):
    """1M database vectors x 64 sub-quantizers (32 MiB of codes): (i) a sample of columns equals the closed form, (ii) the
    result does not depend on the block size, (iii) the sums are additive in the table (LUT a + LUT b, no overflow),
    (iv) a constant table gives nsq * c everywhere."""
    rng = np.random.default_rng(11)
    M, ntotal, nq = 64, 1 << 20, 4
    codes = rng.integers(0, 256, size=(ntotal, M // 2), dtype=np.uint8)
    la = rng.integers(0, 128, size=(nq, M, 16), dtype=np.uint8)
    lb = rng.integers(0, 128, size=(nq, M, 16), dtype=np.uint8)
    dcodes = cu(codes)
    res = {}
    for bbs in (32, 256):
        blocks = pq4.pq4_pack_codes(dcodes, ntotal, M, ntotal, bbs, M)
        res[bbs] = [pq4.pq4_accumulate_loop(nq, ntotal, bbs, M, blocks, pq4.pq4_pack_LUT(nq, M, cu(t))) for t in (la, lb, la + lb)]
    a, b, ab = res[32]
    assert all(torch.equal(x.view(torch.int16), y.view(torch.int16)) for x, y in zip(res[32], res[256]))
    assert torch.equal(a.to(torch.int32) + b.to(torch.int32), ab.to(torch.int32))
    cols = rng.integers(0, ntotal, 4096)
    assert np.array_equal(host(a)[:, cols], orc.np_pq4_distances(codes[cols], la, M))
    const = np.full((1, M, 16), 3, dtype=np.uint8)
    blocks = pq4.pq4_pack_codes(dcodes, ntotal, M, ntotal, 32, M)
    c = pq4.accumulate_to_mem(1, ntotal, M, blocks, pq4.pq4_pack_LUT(1, M, cu(const)))
    c32 = c.to(torch.int32)
    assert int(c32.min()) == int(c32.max()) == 3 * M


def test_pq4_errors(pq4):
    from rvv_lut_vq_b200._lib import LutvqError
    z = lambda n: pq4.aligned_empty(n, "cuda")
    with pytest.raises(LutvqError):                  # nb % bbs (pq4_fast_scan_search_1.cpp:150)
        pq4.pq4_accumulate_loop(1, 32, 64, 2, z(64), z(32))
    with pytest.raises(LutvqError):                  # nq = 5
        pq4.pq4_accumulate_loop(5, 32, 32, 2, z(32), z(5 * 32))
    with pytest.raises(LutvqError):                  # misaligned codes (pq4_fast_scan_search_1.cpp:147)
        pq4.pq4_accumulate_loop(1, 32, 32, 2, z(64)[8:40], z(32))
    with pytest.raises(LutvqError):                  # bbs % 32 (pq4_fast_scan.cpp:53)
        pq4.pq4_pack_codes(z(48), 48, 2, 48, 48, 2)
    with pytest.raises(LutvqError):                  # qbs digit 5
        pq4.pq4_accumulate_loop_qbs(0x15, 32, 2, z(32), z(6 * 32))
    with pytest.raises(RuntimeError):                # CPU tensors: no fallback
        pq4.pq4_pack_LUT(1, 2, torch.zeros(32, dtype=torch.uint8))
    assert pq4.pq4_accumulate_loop(1, 0, 32, 2, z(32), z(32)).shape == (1, 0)      # empty database
