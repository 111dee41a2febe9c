#!/usr/bin/env python
"""tools/check_sharded.py -- run under torchrun on N GPUs: the sharded long-signal path
(halo exchange over NCCL + simd_b200_filter_shard per rank) against the single-GPU result
of the same signal computed on rank 0."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import veles.simd_b200 as v  # noqa: E402
from veles.simd_b200 import sharded  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
v.use_torch_stream()
ok = True
for x_len, hl, rev in [(1_000_003, 8192, False), (300_001, 129, True), (4_000_000, 4096, False)]:
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.rand((x_len,), generator=g, device="cuda") * 2 - 1      # same on every rank
    h = torch.rand((hl,), generator=g, device="cuda") * 2 - 1
    s, e = sharded.signal_range(x_len, rank, world)
    ext, loc = sharded.alloc_shard(e - s, hl - 1, device="cuda")
    loc.copy_(x[s:e])
    part = sharded.filter_long_signal(loc, h, x_len, rank, world, sharded.gpu_compute, reverse=rev, ext=ext)
    torch.cuda.synchronize()
    full = v.convolve(x, h, correlate=rev)
    torch.cuda.synchronize()
    want = full[s:s + part.numel()]
    err = (torch.linalg.vector_norm((part - want).double()) / torch.linalg.vector_norm(want.double())).item()
    flag = torch.tensor([1.0 if err < 2e-6 else 0.0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"x={x_len} h={hl} reverse={rev}: all ranks match = {bool(flag.item())} (rank0 rel-L2 {err:.2e})")
    ok = ok and bool(flag.item())
dist.destroy_process_group()
sys.exit(0 if ok else 1)
