"""Multi-GPU partitioning of the convolve / correlate path (one process per GPU,
torch.distributed for the plumbing; NCCL on GPUs, gloo on CPU for the tests).

The reference has no parallelism at all on this path (/root/reference/src/convolve.c:179-180
declines even OpenMP), so this layer is new design:

  * batches of independent signals: rank r owns rows [r*B/W, (r+1)*B/W); h is replicated;
    every rank runs the single-GPU batched call on its rows.  No data-path collective.
  * one long signal: rank r owns x[s_r, e_r) (contiguous, pre-sharded, device resident) and
    produces result[s_r, e_r) -- the last rank also the h-1 tail samples.  The only
    exchange is the (h-1)-sample halo: each rank sends its last h-1 samples to rank r+1
    (32 KB for h = 8192), point to point.  No all-reduce, no all-gather.

`compute` is the function that filters one shard.  On GPUs it is
veles.simd_b200.filter_shard (the C-ABI shard primitive simd_b200_filter_shard); the CPU
tests inject the oracle there to exercise exactly this partition / halo logic under gloo.
"""
import torch
import torch.distributed as dist


def batch_rows(batch, rank, world):
    """Row range [lo, hi) of `batch` independent signals owned by `rank` (balanced)."""
    base, rem = divmod(batch, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def signal_range(x_length, rank, world):
    """Sample range [s, e) of a long signal owned by `rank` (balanced, contiguous)."""
    return batch_rows(x_length, rank, world)


def alloc_shard(n_local, halo, device=None, dtype=torch.float32):
    """Storage for a shard with room for its halo in front: returns (ext, local) where
    local = ext[halo:] is where the rank keeps its samples and ext[:halo] receives the
    neighbour's tail in place (no concatenation copy per call)."""
    # the shard itself starts on a 16-byte boundary (the kernels' wide-load paths want that);
    # the halo in front of it may then start up to 3 samples off a boundary
    lead = (-halo) % 4
    buf = torch.zeros(lead + halo + n_local, dtype=dtype, device=device)
    ext = buf[lead:]
    return ext, ext[halo:]


def exchange_halo(x_local, halo, rank, world, group=None, into=None):
    """Returns the `halo` samples that precede this rank's shard (zeros on rank 0): the tail
    of rank-1's shard, received point to point while this rank's own tail goes to rank+1.
    `into` (optional) is the tensor that receives them.  Shards shorter than the halo are
    not supported (they would need a multi-hop halo)."""
    assert x_local.dim() == 1
    if halo == 0:
        return x_local.new_zeros(0)
    assert x_local.numel() >= halo, "shard shorter than the h-1 halo"
    left = into if into is not None else x_local.new_zeros(halo)
    ops = []
    if rank + 1 < world:
        ops.append(dist.P2POp(dist.isend, x_local[-halo:].contiguous(), rank + 1, group))
    if rank > 0:
        ops.append(dist.P2POp(dist.irecv, left, rank - 1, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return left


def filter_long_signal(x_local, h, x_length, rank, world, compute, reverse=False, group=None,
                       ext=None, out=None, overlap=False):
    """This rank's part of convolve / cross_correlate of one signal of x_length samples that
    is sharded by signal_range().  Returns result[s_r, e_r) (plus the h-1 tail on the last
    rank), so that concatenating the ranks' returns gives the full x_length+h-1 result.

    compute(x, h, reverse, left_pad, result_length, out=None) -> tensor  is the shard primitive.
    ext: the buffer from alloc_shard() whose tail is x_local; the halo is then received in
    place in front of the shard.  out: optional preallocated result tensor.

    overlap=True computes the outputs that depend only on this rank's own samples (from index
    h-1 on) while the halo travels and the first h-1 outputs afterwards; on NVLink the 32 KB
    exchange is as short as that extra small launch, so the default is the simple order
    (exchange, then one launch).
    """
    hl = h.numel()
    halo = hl - 1
    s, e = signal_range(x_length, rank, world)
    n_local = e - s
    assert x_local.numel() == n_local
    n_out = n_local + (halo if rank == world - 1 else 0)
    if rank == 0:
        # no predecessor: the h-1 samples before x[0] are the implicit zeros of a full
        # convolution (left_pad = h-1); still send our tail to rank 1
        if world > 1 and halo > 0:
            exchange_halo(x_local, halo, rank, world, group)
        return _call(compute, x_local, h, reverse, halo, n_out, out)
    if halo == 0:
        return _call(compute, x_local, h, reverse, 0, n_out, out)
    assert n_local >= halo, "shard shorter than the h-1 halo"
    if out is None:
        out = x_local.new_empty(n_out)
    if ext is None:
        ext = torch.cat([x_local.new_zeros(halo), x_local])
    if not overlap:
        exchange_halo(x_local, halo, rank, world, group, into=ext[:halo])
        return _call(compute, ext, h, reverse, 0, n_out, out)
    # 1. start the exchange (tail of this shard -> rank+1, halo <- rank-1)
    ops = []
    if rank + 1 < world:
        ops.append(dist.P2POp(dist.isend, x_local[-halo:].contiguous(), rank + 1, group))
    ops.append(dist.P2POp(dist.irecv, ext[:halo], rank - 1, group))
    reqs = dist.batch_isend_irecv(ops)
    # 2. everything that needs no halo: result[head:] from x_local alone.  `head` is the halo
    #    rounded up to 4 samples so that the big piece of the result starts 16-byte aligned.
    head = min(-(-halo // 4) * 4, n_out)
    if n_out > head:
        _call(compute, x_local[head - halo:], h, reverse, 0, n_out - head, out[head:])
    # 3. the first `head` outputs, from [halo | first samples of the shard]
    for req in reqs:
        req.wait()
    _call(compute, ext[:head + halo], h, reverse, 0, head, out[:head])
    return out


def _call(compute, x, h, reverse, left_pad, result_length, out):
    if out is None:
        return compute(x, h, reverse, left_pad, result_length)
    res = compute(x, h, reverse, left_pad, result_length, out)
    if res is not out and res is not None:
        out.copy_(res)
    return out


def gpu_compute(x, h, reverse, left_pad, result_length, out=None):
    """The B200 shard primitive (simd_b200_filter_shard through the C-ABI)."""
    import veles.simd_b200 as v
    return v.filter_shard(x, h, reverse, left_pad, result_length, out=out)
