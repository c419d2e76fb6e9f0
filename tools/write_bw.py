#!/usr/bin/env python
"""write-only HBM bandwidth on this GPU (fill of 1 GiB, and of a ring of 90 MB buffers): the ceiling of the dequantize kernel's output stream"""
import json
import torch
dev = torch.device("cuda:0")
big = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
ring = [torch.empty(90_177_536, dtype=torch.uint8, device=dev) for _ in range(12)]
def timeit(fn, n):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / n
us = timeit(lambda: big.fill_(1), 10)
print(json.dumps({"op": "fill 1 GiB", "us": round(us, 1), "write_gbps": round((1 << 30) / us / 1e3, 1)}))
k = [0]
def ringfill():
    ring[k[0] % 12].fill_(1); k[0] += 1
us = timeit(ringfill, 48)
print(json.dumps({"op": "fill 90 MB (ring of 12)", "us": round(us, 1), "write_gbps": round(90_177_536 / us / 1e3, 1)}))
src = torch.empty(1 << 29, dtype=torch.uint8, device=dev); dst = torch.empty(1 << 29, dtype=torch.uint8, device=dev)
us = timeit(lambda: dst.copy_(src), 10)
print(json.dumps({"op": "copy 512 MiB", "us": round(us, 1), "read_plus_write_gbps": round(2 * (1 << 29) / us / 1e3, 1)}))
