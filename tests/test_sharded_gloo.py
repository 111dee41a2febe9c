"""CPU tests (gloo, world_size 2 and 3) of the multi-GPU partitioning logic: batch row
ownership and the long-signal halo exchange.  The shard primitive is the CPU oracle here
(test infrastructure); on GPUs the same code path calls simd_b200_filter_shard."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_compute(x, h, reverse, left_pad, result_length, out_t=None):
    """result[n] = sum_j g[j] X(n + j - left_pad) via the oracle's full convolution."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _oracle import Oracle
    o = Oracle()
    xn, hn = x.numpy(), h.numpy()
    full = o.convolve_f64(xn, hn, int(reverse))       # full[k] = sum_j g[j] X(k + j - (hl-1))
    off = hn.size - 1 - left_pad                      # result[n] = full[n + off]
    out = np.zeros(result_length, dtype=np.float64)
    n = min(result_length, full.size - off)
    out[:n] = full[off:off + n]
    res = torch.from_numpy(out.astype(np.float32))
    if out_t is not None:
        out_t.copy_(res)
        return out_t
    return res


def _worker(rank, world, port, x_length, hl, reverse, q, overlap=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    from veles.simd_b200 import sharded
    rng = np.random.default_rng(123)
    x = torch.from_numpy(rng.uniform(-1, 1, x_length).astype(np.float32))
    h = torch.from_numpy(rng.uniform(-1, 1, hl).astype(np.float32))
    s, e = sharded.signal_range(x_length, rank, world)
    ext, loc = sharded.alloc_shard(e - s, hl - 1)
    assert loc.data_ptr() % 16 == 0
    loc.copy_(x[s:e])
    part = sharded.filter_long_signal(loc, h, x_length, rank, world, _oracle_compute,
                                      reverse=reverse, ext=ext, overlap=overlap)
    gathered = [None] * world
    dist.all_gather_object(gathered, part.numpy())
    if rank == 0:
        q.put(np.concatenate(gathered))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,x_length,hl,reverse,overlap", [(2, 5000, 129, False, False),
                                                              (2, 4001, 700, True, True),
                                                              (3, 9000, 64, False, True)])
def test_long_signal_halo_partition(world, x_length, hl, reverse, overlap):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _oracle import Oracle, rel_l2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, x_length, hl, reverse, q, overlap))
             for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(123)
    x = rng.uniform(-1, 1, x_length).astype(np.float32)
    h = rng.uniform(-1, 1, hl).astype(np.float32)
    want = Oracle().convolve_f64(x, h, int(reverse))
    assert got.shape == want.shape
    assert rel_l2(got, want) < 1e-6


def test_batch_rows_cover_batch_exactly():
    sys.path.insert(0, ROOT)
    from veles.simd_b200 import sharded
    for batch in (1, 7, 1024, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            ranges = [sharded.batch_rows(batch, r, world) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == batch
            for a, b in zip(ranges, ranges[1:]):
                assert a[1] == b[0]
            sizes = [hi - lo for lo, hi in ranges]
            assert max(sizes) - min(sizes) <= 1
