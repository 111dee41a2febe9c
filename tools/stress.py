#!/usr/bin/env python
"""tools/stress.py [seconds] -- randomized parity sweep on the GPU: random shapes through
every entry-point family (direct, overlap-save, FFT, best handle, batch, shard, no-alloc,
real-packed forced, wavelet transform, host and device pointers) against the CPU oracle.  Prints a summary and
exits non-zero at the first mismatch."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import veles.simd_b200 as v  # noqa: E402
from _oracle import Oracle, parity_ok  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
rng = np.random.default_rng(int(time.time()) & 0xFFFF)
o = Oracle()
v.use_torch_stream()
t_end = time.time() + budget
n_cases = 0
worst = 0.0


def want_of(x, h, corr):
    if x.size * h.size <= 4_000_000:
        return o.convolve_f64(x, h, corr)
    return o.convolve_fft(x, h, corr)


def check(got, x, h, corr, what):
    global worst
    ok, r, m = parity_ok(got, want_of(x, h, corr), x, h)
    worst = max(worst, r)
    if not ok:
        print("MISMATCH", what, x.size, h.size, corr, r, m)
        sys.exit(1)


while time.time() < t_end:
    xl = int(2 ** rng.uniform(0, 19))
    hl = int(2 ** rng.uniform(0, min(15, np.log2(xl) + 1.5)))
    xl, hl = max(xl, 1), max(hl, 1)
    corr = bool(rng.integers(0, 2))
    x = rng.uniform(-1, 1, xl).astype(np.float32)
    h = rng.uniform(-1, 1, hl).astype(np.float32)
    dev = bool(rng.integers(0, 2))
    xa = torch.from_numpy(x).cuda() if dev else x
    ha = torch.from_numpy(h).cuda() if dev else h

    def host(a):
        if dev:
            torch.cuda.synchronize()
            return a.cpu().numpy()
        return a

    if xl * hl <= 50_000_000:
        check(host(v.convolve_simd(xa, ha, correlate=corr)), x, h, corr, "simd")
    with v.Handle(xl, hl, "", corr) as hd:
        check(host(hd(xa, ha)), x, h, corr, f"best alg={hd.algorithm}")
    if xl + hl < 300000:
        with v.Handle(xl, hl, "_fft", corr) as hd:
            check(host(hd(xa, ha)), x, h, corr, "fft")
    if hl < xl // 2:
        for real in ("0", "1"):
            os.environ["VELES_B200_REAL"] = real
            with v.Handle(xl, hl, "_overlap_save", corr) as hd:
                check(host(hd(xa, ha)), x, h, corr, f"os real={real} L={hd.transform_length}")
        os.environ.pop("VELES_B200_REAL")
    if hl <= 8192 and xl >= 1:
        xd, hd_ = torch.from_numpy(x).cuda(), torch.from_numpy(h).cuda()
        out = torch.empty(xl + hl - 1, device="cuda")
        rc = v.lib.simd_b200_filter_noalloc(xd.data_ptr(), xl, hd_.data_ptr(), hl, int(corr), out.data_ptr(),
                                            torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        torch.cuda.synchronize()
        check(out.cpu().numpy(), x, h, corr, "noalloc")
    if xl > 4 * hl and xl > 64:
        cut = int(rng.integers(hl, xl - 1))
        full = want_of(x, h, corr)
        a = v.filter_shard(x[:cut], h, int(corr), hl - 1, cut)
        b = v.filter_shard(x[cut - (hl - 1):], h, int(corr), 0, xl + hl - 1 - cut)
        ok, r, m = parity_ok(np.concatenate([a, b]), full, x, h)
        if not ok:
            print("MISMATCH shard", xl, hl, cut, r, m)
            sys.exit(1)
    if xl * 8 < 4_000_000:
        B = int(rng.integers(2, 9))
        xb = rng.uniform(-1, 1, (B, xl)).astype(np.float32)
        xba = torch.from_numpy(xb).cuda() if dev else xb
        with v.Handle(xl, hl, "", corr) as hd:
            yb = host(hd(xba, ha))
        for b in (0, B - 1):
            check(yb[b], xb[b], h, corr, "batch")
    # wavelet transform: random family / order / extension / level on a random even length
    fam = int(rng.integers(0, 3))
    order = int(rng.integers(1, 6)) * 6 if fam == 1 else int(rng.integers(1, 39)) * 2
    ext = int(rng.integers(0, 4))
    wl = max(2, int(2 ** rng.uniform(1, 17)) & ~1)
    level = None if rng.integers(0, 2) else int(rng.integers(1, 7))
    w = rng.uniform(-1, 1, wl).astype(np.float32)
    hp, lp = v.wavelet_filters(fam, order)
    want_hi, want_lo = o.wavelet_apply(lp, ext, w, level)
    wa = torch.from_numpy(w).cuda() if dev else w
    hi, lo = v.wavelet_apply(fam, order, ext, wa, level)
    for got, want, filt in ((host(hi), want_hi, hp), (host(lo), want_lo, lp)):
        scale = float(np.abs(filt).sum())
        err = float(np.abs(got.astype(np.float64) - want).max())
        r = float(np.linalg.norm(got.astype(np.float64) - want) / max(np.linalg.norm(want), 1e-30))
        if not (err <= 1e-4 * scale and (r <= 1e-5 or err <= 2 * order * 2.0 ** -24 * scale)):
            print("MISMATCH wavelet", fam, order, ext, wl, level, r, err)
            sys.exit(1)
    n_cases += 1
print(f"stress OK: {n_cases} random shapes, worst rel-L2 {worst:.2e}")
