"""GPU tests of the boundary work of round 2: the small-call path of the reference API, the
two PCIe-saving chains (int16 samples in, extrema only out) and the one-process multi-GPU
partitioning, all through the C-ABI and against the CPU oracle."""
import os

import numpy as np
import pytest

from _oracle import parity_ok, rel_l2

pytestmark = pytest.mark.gpu


def uni(seed, *shape):
    return np.random.default_rng(seed).uniform(-1, 1, shape).astype(np.float32)


def check(got, want, x, h, what=""):
    ok, r, m = parity_ok(got, want, x, h)
    assert ok, f"{what}: rel_l2={r:.3e} max_abs={m:.3e}"


@pytest.mark.parametrize("xl,hl", [(1024, 50), (8, 4), (1000, 999), (20000, 300), (30000, 4096), (65536 - 600, 512)])
def test_small_host_calls_every_family(simd, oracle, xl, hl):
    """Host-pointer calls below 256 KiB take the one-stream bounce path (direct kernel:
    zero-copy through mapped pinned memory; transform path: one copy each way)."""
    x, h = uni(xl, xl), uni(hl + 1, hl)
    want = oracle.convolve_f64(x, h).astype(np.float32)
    check(simd.convolve_simd(x, h), want, x, h, "simd")
    for kind in ("", "_fft") + (("_overlap_save",) if hl < xl // 2 else ()):
        with simd.Handle(xl, hl, kind) as hd:
            for call in range(2):                      # the bounce buffer is reused
                check(hd(x, h), want, x, h, f"{kind} call {call}")
    wantc = oracle.convolve_f64(x, h, reverse=1).astype(np.float32)
    with simd.Handle(xl, hl, "", True) as hd:
        check(hd(x, h), wantc, x, h, "correlate")


def test_small_host_calls_from_threads(simd, oracle):
    """The bounce buffer is per device and shared: concurrent callers must serialise on it."""
    import threading
    errs = []

    def work(seed):
        try:
            for i in range(20):
                x, h = uni(seed * 100 + i, 1500 + seed), uni(seed * 100 + i + 7, 40 + seed)
                got = simd.convolve_simd(x, h)
                ok, r, m = parity_ok(got, oracle.convolve_brute(x, h), x, h)
                assert ok, (seed, i, r, m)
        except Exception as e:        # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=work, args=(s,)) for s in range(4)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs


@pytest.mark.parametrize("B,xl,hl,corr", [(6, 5000, 33, False), (3, 70000, 700, True), (40, 16384, 64, False)])
def test_int16_samples_are_widened_on_the_device(simd, oracle, B, xl, hl, corr):
    rng = np.random.default_rng(B * xl)
    pitch = xl + 6                                       # strided rows: the 2-D copy path
    xbuf = rng.integers(-32768, 32767, (B, pitch), dtype=np.int16)
    x16 = xbuf[:, :xl]
    h = uni(hl, hl)
    with simd.Handle(xl, hl, "", corr) as hd:
        ptr = xbuf.ctypes.data
        out = np.empty((B, xl + hl - 1), np.float32)
        getattr(simd.lib, hd.prefix + "_batch_int16")(hd.handle, ptr, pitch, h.ctypes.data, out.ctypes.data,
                                                      out.shape[1], B)
    for b in range(B):
        xf = oracle.int16_to_float(np.ascontiguousarray(x16[b]))      # reference converter, restated
        want = oracle.convolve_f64(xf, h, reverse=1 if corr else 0)
        ok, r, m = parity_ok(out[b], want.astype(np.float32), xf, h)
        assert ok, (b, r, m)


def test_correlate_then_peaks_returns_only_the_extrema(simd, oracle):
    xl, hl = 300000, 257
    t = np.arange(xl, dtype=np.float32)
    x = (np.sin(t * 0.01) + 0.3 * np.sin(t * 0.37)).astype(np.float32) + 0.01 * uni(1, xl)
    h = np.hanning(hl).astype(np.float32)
    pos, val = simd.cross_correlate_peaks(x, h)
    # the same chain as two calls of the library ...
    y = simd.cross_correlate(x, h)
    pos2, val2 = simd.detect_peaks(y)
    assert np.array_equal(pos, pos2) and np.array_equal(val, val2)
    # ... and the oracle's scalar detect_peaks on that correlation result
    wpos, wval = oracle.detect_peaks(y)
    assert np.array_equal(pos, wpos) and np.array_equal(val, wval)
    check(y, oracle.convolve_f64(x, h, reverse=1).astype(np.float32), x, h, "correlation itself")
    assert 0 < pos.size < xl // 4


@pytest.fixture
def two_devices(simd, monkeypatch):
    """Two GPUs when the box has them, otherwise two slices sharing the one GPU (test knob)."""
    monkeypatch.setenv("VELES_B200_OVERSUBSCRIBE", "1")
    n = simd.use_devices(3)
    assert n == 3
    yield n
    simd.use_devices(1)


def test_batch_spread_over_devices_from_one_process(simd, oracle, two_devices):
    B, xl, hl = 37, 100000, 300                          # 37 rows over 3 slices: 13 + 12 + 12
    x, h = uni(5, B, xl), uni(6, hl)
    with simd.Handle(xl, hl) as hd:
        y = hd(x, h)
    for b in (0, 12, 13, 24, 25, 36):
        check(y[b], oracle.convolve_f64(x[b], h).astype(np.float32), x[b], h, f"row {b}")
    s = y.astype(np.float64).sum(axis=1)
    assert np.allclose(s, x.astype(np.float64).sum(axis=1) * h.astype(np.float64).sum(), rtol=0, atol=1e-2)


@pytest.mark.parametrize("corr", [False, True])
def test_long_host_signal_spread_over_devices(simd, oracle, two_devices, corr):
    xl, hl = 5_000_001, 1000
    x, h = uni(8, xl), uni(9, hl)
    y = simd.filter_long(x, h, reverse=corr)
    want = oracle.convolve_fft(x, h, reverse=1 if corr else 0)
    check(y, want, x, h, "filter_long")
    # the seams between the three output ranges
    per = -(-(xl + hl - 1) // 3)
    for seam in (per, 2 * per):
        assert rel_l2(y[seam - 2000:seam + 2000], want[seam - 2000:seam + 2000]) < 1e-5


def test_device_resident_shards_with_peer_halo(simd, oracle):
    """simd_b200_filter_long_sharded: shards on the GPUs present (all on one GPU when there is
    only one: the halo copy is then a local copy, the partition logic is the same)."""
    import torch
    from veles.simd_b200 import sharded
    ndev = torch.cuda.device_count()
    n_shards, xl, hl = 4, 3_000_003, 2049
    x, h = uni(11, xl), uni(12, hl)
    exts, locs, hs = [], [], []
    for i in range(n_shards):
        dev = torch.device("cuda", i % ndev)
        s, e = sharded.signal_range(xl, i, n_shards)
        ext, loc = sharded.alloc_shard(e - s, hl - 1, device=dev)
        loc.copy_(torch.from_numpy(x[s:e]))
        exts.append(ext); locs.append(loc); hs.append(torch.from_numpy(h).to(dev))
    for d in range(ndev):
        torch.cuda.synchronize(d)
    res = simd.filter_long_sharded(locs, hs)
    y = np.concatenate([r.cpu().numpy() for r in res])
    assert y.size == xl + hl - 1
    check(y, oracle.convolve_fft(x, h), x, h, "sharded")


def test_pageable_and_pinned_host_buffers_agree(simd, oracle):
    """Pageable caller memory goes through pinned bounce slots filled by copy threads, pinned
    memory is read by the copy engine in place; strided rows exercise the compaction."""
    import torch
    B, xl, hl, xp, op = 48, 90000, 500, 90016, 90600
    xbuf = uni(21, B, xp)
    h = uni(22, hl)
    ybuf = np.full((B, op), np.nan, np.float32)
    x, y = xbuf[:, :xl], ybuf[:, :xl + hl - 1]
    with simd.Handle(xl, hl) as hd:
        simd.lib.convolve_batch(hd.handle, xbuf.ctypes.data, xp, h.ctypes.data, ybuf.ctypes.data, op, B)
        xpin = torch.from_numpy(np.ascontiguousarray(x)).pin_memory()
        ypin = torch.empty((B, xl + hl - 1)).pin_memory()
        hd(xpin, torch.from_numpy(h).pin_memory(), out=ypin)
    assert np.array_equal(y, ypin.numpy())
    assert np.isnan(ybuf[:, xl + hl - 1:]).all()            # nothing written between the rows
    for b in (0, 17, B - 1):
        check(y[b], oracle.convolve_fft(np.ascontiguousarray(x[b]), h), x[b], h, f"row {b}")
