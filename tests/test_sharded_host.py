"""CPU: host logic of the output-channel-sharded path -- row partition, handle exchange and the
slice/all-gather algebra -- with world_size 2 on gloo.  No CUDA call is made here."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_rows_partition():
    from rvv_lut_vq_b200.sharded import shard_rows
    for O in (32, 1000, 1024, 4096, 11008, 28672, 32000, 8192, 33):
        for P in (1, 2, 4, 8):
            prev = 0
            for r in range(P):
                b, e = shard_rows(O, P, r)
                assert b == prev and e >= b
                assert b % 32 == 0 or b == O                 # whole row sets: the shard's packed image needs no re-split
                prev = e
            assert prev == O
            sizes = [shard_rows(O, P, r)[1] - shard_rows(O, P, r)[0] for r in range(P)]
            assert max(sizes) - min(sizes) <= 32 or O < 32 * P
    with pytest.raises(ValueError):
        shard_rows(64, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from rvv_lut_vq_b200 import synth
        from rvv_lut_vq_b200.sharded import HANDLE_BYTES, exchange_handles, shard_rows
        import oracle as orc
        # (1) handle exchange: every rank ends up with every rank's 64 bytes, indexed by rank
        mine = bytes([rank * 16 + (i % 16) for i in range(HANDLE_BYTES)])
        allh = exchange_handles(mine)
        assert len(allh) == world * HANDLE_BYTES
        for r in range(world):
            assert allh[r * HANDLE_BYTES:(r + 1) * HANDLE_BYTES] == bytes([r * 16 + (i % 16) for i in range(HANDLE_BYTES)])
        # (2) the sharded algebra on the oracle: slice of codes/scales/bias -> local GEMV -> all-gather == full GEMV
        I, O = 512, 1000                                     # 1000 rows: ranks own 512 and 488 rows
        L = synth.make_layer_np(I, O, seed=11, scales="randn", bias=True)
        b, e = shard_rows(O, world, rank)
        full = orc.c_aqlm_gemv_rowmajor(L["x"][0], L["codes"], L["codebooks"], L["scales"], L["bias"])
        local = orc.c_aqlm_gemv_rowmajor(L["x"][0], L["codes"][b:e], L["codebooks"], L["scales"].reshape(-1)[b:e], L["bias"][b:e])
        assert local.shape[0] == e - b
        max_rows = max(shard_rows(O, world, r)[1] - shard_rows(O, world, r)[0] for r in range(world))
        pad = np.zeros(max_rows, dtype=np.uint16)
        pad[: e - b] = local
        out = torch.empty(world * max_rows * 2, dtype=torch.uint8)             # gloo moves bytes
        dist.all_gather_into_tensor(out, torch.from_numpy(pad.view(np.uint8)))
        got = np.concatenate([out.numpy().view(np.uint16)[r * max_rows: r * max_rows + (shard_rows(O, world, r)[1] - shard_rows(O, world, r)[0])]
                              for r in range(world)])
        assert np.array_equal(got, full), "gathered slices differ from the unsharded result"
        with open(os.path.join(tmp, f"ok{rank}"), "w") as f:
            f.write("ok")
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo(tmp_path, orc):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}") for r in range(world))
