"""GPU: the sharded decode step through the C-ABI.
 * one GPU: P ranks emulated in one process (lutvq_shard_attach_ptrs) -- the push / epoch / wait protocol and the
   bit-identity of gathered slices with the unsharded GEMV;
 * >= 2 GPUs (skipped on a one-GPU box): one process per GPU, CUDA-IPC attach, push and NCCL modes."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def cu(a, dev="cuda:0"):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


@pytest.mark.parametrize("world,I,O", [(1, 512, 256), (2, 1024, 1000), (4, 4096, 4096), (8, 2048, 1024)])
def test_push_protocol_one_gpu(world, I, O):
    from rvv_lut_vq_b200 import ops, synth
    from rvv_lut_vq_b200.sharded import ShardGroup, ShardedQuantizedLinear, shard_rows
    L = synth.make_layer_np(I, O, seed=world * 7 + 1, scales="randn", bias=True)
    codes, cb, sc, bi = cu(L["codes"]), cu(L["codebooks"]), cu(L["scales"]), cu(L["bias"])
    full = ops.code2x8_matmat(cu(L["x"]), codes, cb, sc, bi)[0]
    groups = [ShardGroup(world, r, O, n_slots=4) for r in range(world)]
    for g in groups:
        g.attach_same_process(groups)
    layers = [ShardedQuantizedLinear(codes, cb, sc, bi, O, world, r, group=groups[r], mode="push") for r in range(world)]
    streams = [torch.cuda.Stream() for _ in range(world)]
    torch.cuda.synchronize()
    for rep in range(6):                                   # more reps than slots: slot reuse + epochs
        x = cu(L["x"])[0] * (1.0 + rep)
        want = ops.code2x8_matmat(x.reshape(1, -1), codes, cb, sc, bi)[0]
        torch.cuda.synchronize()
        slot = rep % 4
        # every "rank" launches its GEMV first (no rank may wait before its peers have pushed: one process, one GPU)
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                g = groups[r]
                lay = layers[r]
                g.matmat(slot, lay.begin, x, lay.packed, lay.codebooks, lay.scales, lay.bias, lay.rows)
        outs = []
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                groups[r].wait(slot)
                outs.append(groups[r].slot_tensor(slot, O, torch.float16).clone())
        torch.cuda.synchronize()
        for r in range(world):
            assert groups[r].error() == 0
            assert torch.equal(outs[r].view(torch.int16), want.view(torch.int16)), f"rank {r} rep {rep}: gathered y differs from the unsharded GEMV"
    assert torch.equal(full.view(torch.int16), ops.code2x8_matmat(cu(L["x"]), codes, cb, sc, bi)[0].view(torch.int16))
    for g in groups:
        g.close()


def test_wait_times_out_instead_of_hanging():
    """a wait with a missing peer returns (bounded spin) and records the error"""
    from rvv_lut_vq_b200.sharded import ShardGroup
    a, b = ShardGroup(2, 0, 64, n_slots=2), ShardGroup(2, 1, 64, n_slots=2)
    a.attach_same_process([a, b])
    b.attach_same_process([a, b])
    a.wait(0)                                              # nobody pushed: times out after ~2 s
    torch.cuda.synchronize()
    assert a.error() != 0
    a.close()
    b.close()


# ------------------------------------------------------------------------------------------------ multi-GPU
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, tmp):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from rvv_lut_vq_b200 import ops, synth
        from rvv_lut_vq_b200.sharded import ShardGroup, ShardedQuantizedLinear
        I, O = 4096, 11008
        L = synth.make_layer_np(I, O, seed=3, scales="randn", bias=True)
        c = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        codes, cb, sc, bi = c(L["codes"]), c(L["codebooks"]), c(L["scales"]), c(L["bias"])
        group = ShardGroup.from_process_group(O, n_slots=4)
        push = ShardedQuantizedLinear(codes, cb, sc, bi, O, world, rank, group=group, mode="push")
        nccl = ShardedQuantizedLinear(codes, cb, sc, bi, O, world, rank, mode="nccl")
        for rep in range(10):
            x = c(L["x"])[0] * (1.0 + 0.5 * rep)
            want = ops.code2x8_matmat(x.reshape(1, -1), codes, cb, sc, bi)[0]
            y1 = push(x).clone()
            y2 = nccl(x).clone()
            torch.cuda.synchronize()
            assert group.error() == 0
            assert torch.equal(y1.view(torch.int16), want.view(torch.int16)), f"push mode differs (rank {rank}, rep {rep})"
            assert torch.equal(y2.view(torch.int16), want.view(torch.int16)), f"nccl mode differs (rank {rank}, rep {rep})"
            dist.barrier()
        group.close()
        with open(os.path.join(tmp, f"ok{rank}"), "w") as f:
            f.write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs (gpurun --gpus 2)")
def test_multi_gpu_push_and_nccl(tmp_path):
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 8)
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}") for r in range(world))
