"""GPU parity tests: the CUDA path, called through the C-ABI (ctypes -> libSimd.so), against
the CPU oracle, the committed golden fixtures (reference outputs) and size-independent
properties at the BASELINE.json sizes.

Tolerance (BASELINE.json north_star): rel-L2 <= 1e-5 and max|err| <= 1e-4*||h||_1*max|x|.
"""
import numpy as np
import pytest

from _oracle import parity_ok, rel_l2

pytestmark = pytest.mark.gpu

KAT_X = np.arange(1, 9, dtype=np.float32)
KAT_H = np.array([10, 9, 8, 7], dtype=np.float32)
KAT_CONV = [10, 29, 56, 90, 124, 158, 192, 226, 170, 113, 56]
KAT_CORR = [7, 22, 46, 80, 114, 148, 182, 216, 187, 142, 80]


def uni(seed, *shape):
    return np.random.default_rng(seed).uniform(-1, 1, shape).astype(np.float32)


def check(got, want, x, h, what=""):
    ok, r, m = parity_ok(got, want, x, h)
    assert ok, f"{what}: rel_l2={r:.3e} max_abs={m:.3e}"
    return r


def test_known_answers_every_entry_point(simd):
    """tests/convolve.cc:53-71 and tests/correlate.cc:53-71 through all call families."""
    for corr, want in ((False, KAT_CONV), (True, KAT_CORR)):
        for s in (0, 1):
            assert np.allclose(simd.convolve_simd(KAT_X, KAT_H, correlate=corr, simd=s), want, atol=1e-4)
        for kind in ("", "_fft"):
            with simd.Handle(8, 4, kind, corr) as hd:
                assert np.allclose(hd(KAT_X, KAT_H), want, atol=1e-4), (kind, corr)
    # overlap-save needs h < x/2 (src/convolve.c:105)
    x = np.concatenate([KAT_X, np.zeros(8, np.float32)])
    for corr, want in ((False, KAT_CONV), (True, KAT_CORR)):
        with simd.Handle(16, 4, "_overlap_save", corr) as hd:
            y = hd(x, KAT_H)
        if corr:
            assert np.allclose(y[:11], want, atol=1e-4)
        else:
            assert np.allclose(y[:11], want, atol=1e-4)


def test_golden_fixtures(simd, golden):
    """Every fixture (outputs of the reference's own code) through the matching entry point."""
    worst = 0.0
    for name, c in golden.items():
        x, h, y = c["x"], c["h"], c["y"]
        corr = c["op"] == "correlate"
        if c["path"] in ("simd0", "simd1"):
            got = simd.convolve_simd(x, h, correlate=corr, simd=int(c["path"][-1]))
        else:
            kind = {"fft": "_fft", "overlap_save": "_overlap_save", "auto": ""}[c["path"]]
            with simd.Handle(x.size, h.size, kind, corr) as hd:
                got = hd(x, h)
                if kind == "":
                    assert hd.algorithm == simd.select_algorithm(x.size, h.size)
        worst = max(worst, check(got, y, x, h, name))
    assert worst < 1e-5


def test_reference_test_shapes(simd, golden):
    """tests/convolve.cc:73-166 / tests/correlate.cc:73-152 with their own criteria."""
    for op in ("convolve", "correlate"):
        corr = op == "correlate"
        for xl, kind in ((1020, "_fft"), (1021, "_overlap_save"), (1024, None)):
            c = golden[f"sin{xl}_{op}_simd0"]
            if kind is None:
                got = simd.convolve_simd(c["x"], c["h"], correlate=corr)
            else:
                with simd.Handle(xl, 50, kind, corr) as hd:
                    got = hd(c["x"], c["h"])
            # sinf(i)*100 inputs: outputs ~1e3, float32 ulp there ~6e-5
            assert np.max((got - c["y"]) ** 2) <= 1e-6 or np.max(np.abs(got - c["y"])) <= 1e-3


@pytest.mark.parametrize("xl,hl", [(1, 1), (2, 1), (1, 7), (7, 7), (33, 5), (100, 100), (257, 17),
                                   (2049, 16), (2048, 48), (5000, 63), (4097, 129), (70000, 64),
                                   (3000, 1500)])
def test_direct_kernel_shapes(simd, oracle, xl, hl):
    x, h = uni(xl, xl), uni(hl + 7, hl)
    for corr in (False, True):
        got = simd.convolve_simd(x, h, correlate=corr)
        check(got, oracle.convolve_f64(x, h, corr), x, h, f"direct {xl}x{hl} corr={corr}")


@pytest.mark.parametrize("hl", [2, 3, 4, 5, 6, 15, 16, 17, 18, 19, 20, 31, 33, 47, 1023, 1024, 1025, 1030,
                                2100])
def test_direct_kernel_tap_tails(simd, oracle, hl):
    """Every remainder of the tap loop (it runs in groups of four taps, sixteen per block) and
    the multi-pass kernel above 1024 taps, at tile-boundary-straddling signal lengths."""
    for xl in (2040, 4100):
        x, h = uni(xl + hl, xl), uni(hl, hl)
        for corr in (False, True):
            got = simd.convolve_simd(x, h, correlate=corr)
            check(got, oracle.convolve_f64(x, h, corr), x, h, f"direct {xl}x{hl} corr={corr}")


@pytest.mark.parametrize("hl", [1, 2, 3, 5, 15, 16, 17, 20, 33, 47, 64, 1023, 1025, 2100])
def test_direct_pipelined_kernel_tap_tails(simd, oracle, hl):
    """The same remainders on a signal long enough (>= 148 tiles of 2048 outputs) that the
    pipelined big-tile kernel runs rather than the small-problem one."""
    xl = 310001
    x, h = uni(xl + hl, xl), uni(hl, hl)
    for corr in (False, True):
        got = simd.convolve_simd(x, h, correlate=corr)
        check(got, oracle.convolve_fft(x, h, corr), x, h, f"direct (pipelined) {xl}x{hl} corr={corr}")


@pytest.mark.parametrize("xl,hl", [(16, 4), (201, 50), (300, 128), (1021, 50), (4099, 129), (9000, 512),
                                   (65536, 512), (20000, 1000), (50000, 2048), (100000, 4096),
                                   (70000, 8192), (100000, 20000), (40000, 12000)])
def test_overlap_save_shapes(simd, oracle, xl, hl):
    """convolve_overlap_save over every transform length, incl. partitioned impulse
    responses (h > 8192) and SURVEY Q1 (300,128): samples past x are zeros."""
    x, h = uni(xl + 1, xl), uni(hl + 2, hl)
    for corr in (False, True):
        with simd.Handle(xl, hl, "_overlap_save", corr) as hd:
            assert hd.reverse == int(corr)
            L = hd.transform_length
            assert 256 <= L <= 16384
            got = hd(x, h)
        want = oracle.convolve_fft(x, h, corr) if hl > 3000 else oracle.convolve_f64(x, h, corr)
        check(got, want, x, h, f"os {xl}x{hl} L={L} corr={corr}")


@pytest.mark.parametrize("xl,hl", [(8, 4), (351, 350), (512, 512), (1020, 50), (5000, 3000),
                                   (9000, 7000), (30000, 30000), (100, 900), (20000, 1000)])
def test_fft_shapes_and_handle_reuse(simd, oracle, xl, hl):
    """convolve_fft: whole-signal transform when x+h-1 <= 16384, partitioned beyond.  The
    handle is reusable (the reference's is single-use)."""
    x, h = uni(xl + 3, xl), uni(hl + 4, hl)
    for corr in (False, True):
        want = oracle.convolve_fft(x, h, corr)
        with simd.Handle(xl, hl, "_fft", corr) as hd:
            M = hd.transform_length
            assert M >= xl + hl - 1 and M & (M - 1) == 0 and M < 2 * (xl + hl - 1)
            for call in range(3):
                check(hd(x, h), want, x, h, f"fft {xl}x{hl} call {call} corr={corr}")


def test_best_method_handle_matches_every_algorithm(simd, oracle):
    for xl, hl in [(1024, 50), (65536, 512), (16384, 64), (6000, 5000), (40000, 300),
                   (1 << 20, 700)]:
        x, h = uni(xl, xl), uni(hl, hl)
        want = oracle.convolve_f64(x, h)
        with simd.Handle(xl, hl) as hd:
            assert (hd.handle.x_length, hd.handle.h_length) == (xl, hl)
            check(hd(x, h), want, x, h, f"auto {xl}x{hl} alg={hd.algorithm}")
        check(simd.convolve_simd(x, h), want, x, h, "direct")
        if hl < xl // 2:
            with simd.Handle(xl, hl, "_overlap_save") as hd:
                check(hd(x, h), want, x, h, "os")
        with simd.Handle(xl, hl, "_fft") as hd:
            check(hd(x, h), want, x, h, "fft")


def test_device_pointers_match_host_pointers(simd):
    import torch
    simd.use_torch_stream()
    for xl, hl in [(1024, 50), (65536, 512), (6000, 5000)]:
        x, h = uni(xl, xl), uni(hl, hl)
        xd, hd_ = torch.from_numpy(x).cuda(), torch.from_numpy(h).cuda()
        for corr in (False, True):
            with simd.Handle(xl, hl, "", corr) as hd:
                host = hd(x, h)
                dev = hd(xd, hd_)
                torch.cuda.synchronize()
                assert np.array_equal(dev.cpu().numpy(), host)
            d2 = simd.convolve_simd(xd, hd_, correlate=corr)
            torch.cuda.synchronize()
            assert np.array_equal(d2.cpu().numpy(), simd.convolve_simd(x, h, correlate=corr))


def test_linearity_and_shift(simd):
    """Properties that hold at any size: conv(a*x1 + x2) = a*conv(x1) + conv(x2); delaying x
    delays the output."""
    xl, hl = 50000, 700
    x1, x2, h = uni(1, xl), uni(2, xl), uni(3, hl)
    with simd.Handle(xl, hl) as hd:
        y1, y2 = hd(x1, h), hd(x2, h)
        y3 = hd((2.5 * x1 + x2).astype(np.float32), h)
    assert rel_l2(y3, 2.5 * y1 + y2) < 5e-6
    xs = np.concatenate([np.zeros(123, np.float32), x1[:-123]])
    with simd.Handle(xl, hl) as hd:
        ys = hd(xs, h)
    # outputs that do not see the 123 samples pushed off the end of x
    assert rel_l2(ys[123:xl], y1[:xl - 123]) < 5e-6


def test_batch_c3_shape(simd, oracle):
    """BASELINE config 3 at full size: 4096 signals x 16384, h = 64 (direct-kernel regime)."""
    import torch
    simd.use_torch_stream()
    B, xl, hl = 4096, 16384, 64
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + 3)
    x = torch.rand((B, xl), generator=g, device="cuda") * 2 - 1
    h = torch.rand((hl,), generator=g, device="cuda") * 2 - 1
    with simd.Handle(xl, hl) as hd:
        assert hd.algorithm == simd.ALG_BRUTE_FORCE
        y = hd(x, h)
        torch.cuda.synchronize()
    assert y.shape == (B, xl + hl - 1)
    hn = h.cpu().numpy()
    for b in (0, 1, 777, 2048, 4095):
        xb = x[b].cpu().numpy()
        check(y[b].cpu().numpy(), oracle.convolve_brute(xb, hn), xb, hn, f"c3 row {b}")
    # sum of a full convolution = (sum x)(sum h), every row (catches any missing tile)
    lhs = y.double().sum(dim=1)
    rhs = x.double().sum(dim=1) * h.double().sum()
    assert torch.max(torch.abs(lhs - rhs)).item() < 1e-2
    # same batch through the correlate entry point equals convolution with flipped h
    with simd.Handle(xl, hl, "", True) as hd:
        yc = hd(x[:64].contiguous(), torch.flip(h, [0]).contiguous())
        torch.cuda.synchronize()
    # (64 signals stay on the direct kernel, 4096 cross over to the transform path)
    assert rel_l2(yc.cpu().numpy(), y[:64].cpu().numpy()) < 2e-6


def test_batch_c4_shape(simd, oracle, reference_or_none):
    """BASELINE config 4 at full size: 1024 signals x 262144, h = 4096 (overlap-save)."""
    import torch
    simd.use_torch_stream()
    B, xl, hl = 1024, 262144, 4096
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + 4)
    x = torch.rand((B, xl), generator=g, device="cuda") * 2 - 1
    h = torch.rand((hl,), generator=g, device="cuda") * 2 - 1
    with simd.Handle(xl, hl) as hd:
        assert hd.algorithm == simd.ALG_OVERLAP_SAVE
        y = hd(x, h)
        torch.cuda.synchronize()
    hn = h.cpu().numpy()
    for b in (0, 511, 1023):
        xb = x[b].cpu().numpy()
        want = oracle.convolve_overlap_save(xb, hn)       # reference algorithm, L = 16384
        check(y[b].cpu().numpy(), want, xb, hn, f"c4 row {b}")
        if reference_or_none is not None and b == 0:
            check(y[b].cpu().numpy(), reference_or_none.handle_call("", xb, hn), xb, hn, "c4 vs _ref")
    # sum of a full convolution = (sum x)(sum h), every row: a dropped or misplaced segment
    # (12289 outputs of magnitude ~20) would move a row sum by ~2e3
    lhs = y.double().sum(dim=1)
    rhs = x.double().sum(dim=1) * h.double().sum()
    tol = 2e-6 * y.double().abs().sum(dim=1)
    assert torch.all(torch.abs(lhs - rhs) <= tol), (torch.max(torch.abs(lhs - rhs)).item(), tol.min().item())
    del y, x
    torch.cuda.empty_cache()


@pytest.mark.parametrize("B,xl,hl,x_pitch,o_pitch", [
    (3, 40000, 4096, 40000, 44095),      # 4 segments: last pair's B partial; odd result pitch -> every store phase
    (2, 30001, 4096, 30004, 34096),      # 3 segments (no B in the last pair), x_len not a multiple of 4
    (5, 13000, 3000, 13000, 16001),      # one pair per signal, A complete, B a handful of outputs
    (2, 12292, 4097, 12292, 16390),      # h-1 a multiple of 4: no extra leading samples; B = 6 outputs (< 16-byte head)
    (1, 200000, 2500, 200000, 202499),   # many pairs of one signal
    (4, 9000, 4096, 9000, 13096),        # segment A already sticks out of the signal
])
def test_persistent_kernel_signal_edges(simd, oracle, monkeypatch, B, xl, hl, x_pitch, o_pitch):
    """N = 16384 paired plan (the persistent Tensor-Memory kernel): first / last pairs of a signal,
    whose parked input range is partly zeros and whose result segments are partial, for every
    16-byte phase of the result rows."""
    import torch
    simd.use_torch_stream()
    monkeypatch.setenv("VELES_B200_FORCE_N", "16384")
    monkeypatch.setenv("VELES_B200_REAL", "0")
    g = torch.Generator(device="cuda").manual_seed(B * 1000 + hl)
    xbuf = torch.rand((B, x_pitch), generator=g, device="cuda") * 2 - 1
    x = xbuf[:, :xl]
    h = torch.rand((hl,), generator=g, device="cuda") * 2 - 1
    ybuf = torch.full((B, o_pitch), float("nan"), device="cuda")
    y = ybuf[:, :xl + hl - 1]
    hd = simd.ConvolutionHandle()
    with simd.Handle(xl, hl, "_overlap_save") as inner:
        assert inner.transform_length == 16384
        hd.algorithm, hd.x_length, hd.h_length = simd.ALG_OVERLAP_SAVE, xl, hl
        hd.handle.os = inner.handle
        simd.lib.convolve_batch(hd, x.data_ptr(), x_pitch, h.data_ptr(), y.data_ptr(), o_pitch, B)
        torch.cuda.synchronize()
    hn = h.cpu().numpy()
    for b in range(B):
        xb = x[b].cpu().numpy().copy()
        check(y[b].cpu().numpy(), oracle.convolve_fft(xb, hn), xb, hn, f"row {b}")
    # nothing written past a row's end
    if o_pitch > xl + hl - 1:
        assert torch.isnan(ybuf[:, xl + hl - 1:]).all()


@pytest.mark.parametrize("seed", range(24))
def test_random_batched_shapes(simd, oracle, seed):
    """Seeded random batches through convolve_batch / cross_correlate_batch on pitched device buffers
    whose rows start at every 4-byte phase (the direct kernel stages 16-byte units only when rows are
    16-byte aligned; the transform kernels pick bulk or scalar stores by the same test), forced onto
    the direct kernel and onto overlap-save in turn.  Row padding must stay untouched."""
    import torch
    simd.use_torch_stream()
    rng = np.random.default_rng(9000 + seed)
    B = int(rng.integers(1, 40))
    hl = int(rng.choice([1, 2, 3, 7, 16, 31, 64, 100, 257, 700]))
    xl = int(rng.integers(max(2 * hl + 2, 8), 60000 if seed % 3 else 400000))
    ol = xl + hl - 1
    x_pitch = xl + int(rng.integers(0, 9))
    o_pitch = ol + int(rng.integers(0, 9))
    x_off, o_off = int(rng.integers(0, 4)), int(rng.integers(0, 4))
    corr = bool(seed & 1)
    g = torch.Generator(device="cuda").manual_seed(seed)
    xbuf = torch.rand((B * x_pitch + 4,), generator=g, device="cuda") * 2 - 1
    h = torch.rand((hl,), generator=g, device="cuda") * 2 - 1
    hn = h.cpu().numpy()
    xrows = xbuf[x_off:x_off + B * x_pitch].view(B, x_pitch)
    call = simd.lib.cross_correlate_batch if corr else simd.lib.convolve_batch
    rows = sorted(set([0, B - 1, B // 2]))
    want = {b: oracle.convolve_fft(xrows[b, :xl].cpu().numpy().copy(), hn, corr) for b in rows}
    for alg in ("brute", "os"):
        ybuf = torch.full((B * o_pitch + 4,), float("nan"), device="cuda")
        yrows = ybuf[o_off:o_off + B * o_pitch].view(B, o_pitch)
        hd = simd.ConvolutionHandle()
        hd.x_length, hd.h_length = xl, hl
        if alg == "brute":
            hd.algorithm = simd.ALG_BRUTE_FORCE
            call(hd, xrows.data_ptr(), x_pitch, h.data_ptr(), yrows.data_ptr(), o_pitch, B)
        else:
            with simd.Handle(xl, hl, "_overlap_save", corr) as inner:
                hd.algorithm = simd.ALG_OVERLAP_SAVE
                hd.handle.os = inner.handle
                call(hd, xrows.data_ptr(), x_pitch, h.data_ptr(), yrows.data_ptr(), o_pitch, B)
                torch.cuda.synchronize()
        torch.cuda.synchronize()
        for b in rows:
            xb = xrows[b, :xl].cpu().numpy().copy()
            check(yrows[b, :ol].cpu().numpy(), want[b], xb, hn,
                  f"seed {seed} {alg} B={B} x={xl} h={hl} corr={corr} row {b}")
        assert not torch.isnan(yrows[:, :ol]).any()
        if o_pitch > ol:
            assert torch.isnan(yrows[:, ol:]).all()
        assert torch.isnan(ybuf[:o_off]).all() and torch.isnan(ybuf[o_off + B * o_pitch:]).all()


def test_long_signal_c5_shape(simd, oracle):
    """BASELINE config 5 at full size: x = 2^30 samples, h = 8192, on one GPU.  Windows of
    the result are checked against a float64 direct convolution of the matching x window."""
    import torch
    simd.use_torch_stream()
    xl, hl = 2 ** 30, 8192
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + 5)
    x = torch.rand((xl,), generator=g, device="cuda") * 2 - 1
    h = torch.rand((hl,), generator=g, device="cuda") * 2 - 1
    with simd.Handle(xl, hl) as hd:
        assert hd.algorithm == simd.ALG_OVERLAP_SAVE
        L = hd.transform_length
        y = hd(x, h)
        torch.cuda.synchronize()
    step = L - hl + 1
    hn = h.cpu().numpy()
    W = 3000
    starts = [0, step - 1500, 7 * step - 10, xl // 2 + 12345, xl - 1500, xl + hl - 1 - W]
    for n0 in starts:
        lo = max(0, n0 - (hl - 1))
        hi = min(xl, n0 + W)
        xw = x[lo:hi].cpu().numpy()
        full = oracle.convolve_f64(xw, hn)             # index k <-> output lo + k
        want = full[n0 - lo:n0 - lo + W]
        got = y[n0:n0 + W].cpu().numpy()
        # windows at the far end include the tail where x has run out: pad is implicit
        ok, r, m = parity_ok(got, want[:got.size], xw, hn)
        assert ok, (n0, r, m)
    total = y.double().sum().item()
    want_total = x.double().sum().item() * h.double().sum().item()
    assert abs(total - want_total) <= 1e-6 * (abs(want_total) + xl ** 0.5 * hl ** 0.5)
    del y, x
    torch.cuda.empty_cache()


def test_filter_shard_reassembles_long_signal(simd, oracle):
    """Multi-GPU partitioning primitive on one GPU: two shards with an (h-1) halo equal the
    whole."""
    xl, hl = 300000, 1000
    x, h = uni(5, xl), uni(6, hl)
    want = oracle.convolve_fft(x, h, 0)
    cut = 150000
    first = simd.filter_shard(x[:cut], h, 0, hl - 1, cut)
    second = simd.filter_shard(x[cut - (hl - 1):], h, 0, 0, xl + hl - 1 - cut)
    got = np.concatenate([first, second])
    check(got, want, x, h, "two shards")
    # forced direct kernel on a shard gives the same numbers to rounding
    d = simd.filter_shard(x[cut - (hl - 1):], h, 0, 0, 5000, algorithm=0)
    assert rel_l2(d, second[:5000]) < 5e-6


def test_host_pointer_pipeline_batches_and_long_signals(simd, oracle):
    """Host buffers flow through the two-slot H2D / kernel / D2H pipeline: a batch is cut
    into groups of signals, one long signal into time slices with halo.  Results must equal
    the device-pointer path bit for bit (same kernels) and the oracle within tolerance."""
    import torch
    simd.use_torch_stream()
    for B, xl, hl in [(40, 70000, 64), (24, 300000, 3000)]:
        x, h = uni(B, B, xl), uni(hl, hl)
        with simd.Handle(xl, hl) as hd:
            host = hd(x, h)
            dev = hd(torch.from_numpy(x).cuda(), torch.from_numpy(h).cuda())
            torch.cuda.synchronize()
        assert np.array_equal(host, dev.cpu().numpy())
        for b in (0, B - 1):
            check(host[b], oracle.convolve_fft(x[b], h, 0), x[b], h, f"pipeline batch row {b}")
    for xl, hl, corr in [(20_000_000, 100, False), (18_000_000, 5000, True)]:
        x, h = uni(xl % 1000, xl), uni(hl, hl)
        with simd.Handle(xl, hl, "", corr) as hd:
            host = hd(x, h)
            dev = hd(torch.from_numpy(x).cuda(), torch.from_numpy(h).cuda())
            torch.cuda.synchronize()
        d = dev.cpu().numpy()
        assert rel_l2(host, d) < 2e-6      # slices may pick another transform length
        for n0 in (0, 8 * 2 ** 20 - 2000, xl - 3000):
            lo = max(0, n0 - (hl - 1))
            want = oracle.convolve_f64(x[lo:n0 + 4000], h, corr)[n0 - lo:n0 - lo + 4000]
            got = host[n0:n0 + 4000]
            ok, r, m = parity_ok(got, want[:got.size], x[lo:n0 + 4000], h)
            assert ok, (xl, hl, n0, r, m)


def test_real_packed_mode(simd, oracle, monkeypatch):
    """The real-packed transform mode (one real segment of 2N samples per transform, used
    when the impulse response is long relative to N) forced on for shapes the cost model
    would not necessarily send there, against the oracle and against the paired mode."""
    import torch
    simd.use_torch_stream()
    for xl, hl, n in [(3000, 100, 256), (4001, 301, 512), (20000, 1000, 1024), (20001, 1500, 2048),
                      (60000, 3000, 4096), (50000, 6000, 8192), (70000, 8192, 16384),
                      (100000, 12000, 16384), (40001, 16384, 16384)]:
        x, h = uni(xl + 11, xl), uni(hl + 12, hl)
        for corr in (False, True):
            want = oracle.convolve_fft(x, h, corr)
            monkeypatch.setenv("VELES_B200_REAL", "1")
            monkeypatch.setenv("VELES_B200_FORCE_N", str(n))
            with simd.Handle(xl, hl, "_overlap_save", corr) as hd:
                assert hd.transform_length == n
                got = hd(x, h)
                got2 = hd(x, h)                      # handle reuse
            monkeypatch.delenv("VELES_B200_REAL")
            monkeypatch.delenv("VELES_B200_FORCE_N")
            check(got, want, x, h, f"real-packed {xl}x{hl} N={n} corr={corr}")
            assert np.array_equal(got, got2)
    # batch + device pointers + shard primitive (pad = 0, odd/even result lengths)
    xl, hl = 90000, 8000
    xb, h = uni(3, 5, xl), uni(4, hl)
    monkeypatch.setenv("VELES_B200_REAL", "1")
    with simd.Handle(xl, hl, "_overlap_save") as hd:
        inner = simd.ConvolutionHandle()
        inner.algorithm, inner.x_length, inner.h_length = 2, xl, hl
        inner.handle.os = hd.handle
        xd, hd_ = torch.from_numpy(xb).cuda(), torch.from_numpy(h).cuda()
        yd = torch.empty((5, xl + hl - 1), device="cuda")
        simd.lib.convolve_batch(inner, xd.data_ptr(), xl, hd_.data_ptr(), yd.data_ptr(), xl + hl - 1, 5)
        torch.cuda.synchronize()
    for b in (0, 4):
        check(yd[b].cpu().numpy(), oracle.convolve_fft(xb[b], h, 0), xb[b], h, f"real-packed batch row {b}")
    x = xb[0]
    full = oracle.convolve_fft(x, h, 0)
    cut = 45001
    second = simd.filter_shard(x[cut - (hl - 1):], h, 0, 0, xl + hl - 1 - cut, algorithm=2)
    check(second, full[cut:], x, h, "real-packed shard")
    monkeypatch.delenv("VELES_B200_REAL")


def test_real_packed_many_segments(simd, oracle, monkeypatch):
    """Real-packed 16384-point transforms over hundreds of segments: odd and even tap counts,
    a last partial segment, batches with a row pitch that is and is not a multiple of four
    samples, and the shard primitive with either parity of pad against h - 1."""
    import torch
    simd.use_torch_stream()
    monkeypatch.setenv("VELES_B200_REAL", "1")
    monkeypatch.setenv("VELES_B200_FORCE_N", "16384")
    try:
        for xl, hl in [(15_000_001, 8192), (16_000_000, 8191), (17_000_003, 6001), (12_500_000, 12000),
                       (10_000_000, 16384)]:
            x, h = uni(xl % 1000, xl), uni(hl, hl)
            for corr in (False, True):
                with simd.Handle(xl, hl, "_overlap_save", corr) as hd:
                    assert hd.transform_length == 16384
                    xd, hd_ = torch.from_numpy(x).cuda(), torch.from_numpy(h).cuda()
                    got = hd(xd, hd_)
                    torch.cuda.synchronize()
                check(got.cpu().numpy(), oracle.convolve_fft(x, h, corr), x, h,
                      f"real-packed long {xl}x{hl} corr={corr}")
        # batch: 40 signals x 30 segments; pitch 600000 (multiple of 4) and 600001 (not)
        hl = 8000
        h = uni(3, hl)
        for xl in (600000, 600001):
            xb = uni(xl, 40, xl)
            with simd.Handle(xl, hl, "") as hd:
                assert hd.algorithm == 2 and hd.transform_length == 16384
                yb = hd(torch.from_numpy(xb).cuda(), torch.from_numpy(h).cuda())
                torch.cuda.synchronize()
            yb = yb.cpu().numpy()
            for b in (0, 17, 39):
                check(yb[b], oracle.convolve_fft(xb[b], h, 0), xb[b], h, f"real-packed batch pitch {xl} row {b}")
        # shard primitive: pad = 0 with an even tap count -> pad and h-1 differ in parity
        x = uni(9, 16_000_000)
        for hl in (8000, 8001):
            h = uni(hl, hl)
            full = oracle.convolve_fft(x, h, 0)
            got = simd.filter_shard(torch.from_numpy(x).cuda(), torch.from_numpy(h).cuda(), 0, 0,
                                    x.size - hl + 1, algorithm=2)
            torch.cuda.synchronize()
            check(got.cpu().numpy(), full[hl - 1:x.size], x, h, f"real-packed shard h={hl}")
    finally:
        monkeypatch.delenv("VELES_B200_REAL")
        monkeypatch.delenv("VELES_B200_FORCE_N")


def test_detect_peaks_parity(simd, oracle):
    """detect_peaks through the C-ABI: bit-exact positions and values against the golden
    fixtures (reference scalar loop) and the oracle, host and device input, all three types,
    sizes around tile boundaries, and the device-resident variant."""
    import os
    import torch
    simd.use_torch_stream()
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "peaks.npz"))
    for name in sorted({k.split("/")[0] for k in z.files}):
        data = z[name + "/data"]
        for t in (1, 2, 3):
            pos, val = simd.detect_peaks(data, t)
            assert np.array_equal(pos, z[f"{name}/pos{t}"]), (name, t)
            assert np.array_equal(val, z[f"{name}/val{t}"]), (name, t)
    rng = np.random.default_rng(8)
    for n in (3, 4, 2049, 2050, 2051, 4098, 100003, 3_000_001):
        data = rng.standard_normal(n).astype(np.float32)
        data[rng.integers(0, n, n // 7)] = 0.0          # plateaus: strict comparisons matter
        for t in (1, 3):
            want_p, want_v = oracle.detect_peaks(data, t)
            got_p, got_v = simd.detect_peaks(data, t)
            assert np.array_equal(got_p, want_p) and np.array_equal(got_v, want_v), (n, t)
            dp, dv = simd.detect_peaks(torch.from_numpy(data).cuda(), t)
            assert np.array_equal(dp, want_p) and np.array_equal(dv, want_v)
    # smooth series: most tiles hold no extremum and are skipped by the write pass; one with none
    t_ = np.arange(2_000_000, dtype=np.float64)
    for data in ((np.sin(t_ * 2e-4) + 0.3 * np.sin(t_ * 7.1e-5)).astype(np.float32),
                 np.arange(500_000, dtype=np.float32)):
        for t in (1, 2, 3):
            want_p, want_v = oracle.detect_peaks(data, t)
            got_p, got_v = simd.detect_peaks(torch.from_numpy(data).cuda(), t)
            assert np.array_equal(got_p, want_p) and np.array_equal(got_v, want_v), t
    # device-resident variant, including truncation at capacity
    data = torch.from_numpy(rng.standard_normal(1_000_000).astype(np.float32)).cuda()
    want_p, want_v = oracle.detect_peaks(data.cpu().numpy(), 3)
    out = torch.zeros((want_p.size + 10, 2), dtype=torch.int32, device="cuda")
    n = simd.detect_peaks_device(data, out, 3)
    assert n == want_p.size
    got = out[:n].cpu().numpy()
    assert np.array_equal(got[:, 0], want_p) and np.array_equal(got[:, 1].copy().view(np.float32), want_v)
    small = torch.zeros((100, 2), dtype=torch.int32, device="cuda")
    assert simd.detect_peaks_device(data, small, 3) == want_p.size
    assert np.array_equal(small.cpu().numpy()[:, 0], want_p[:100])


def test_correlate_then_detect_peaks_on_device(simd, oracle):
    """The consumer chain of SURVEY 8(f): cross-correlate, then pick the extrema of the
    result without the result ever leaving the GPU."""
    import torch
    simd.use_torch_stream()
    xl, hl = 200000, 300
    x, h = uni(31, xl), uni(32, hl)
    xd, hd_ = torch.from_numpy(x).cuda(), torch.from_numpy(h).cuda()
    with simd.Handle(xl, hl, "", True) as hd:
        y = hd(xd, hd_)
        out = torch.zeros((xl + hl, 2), dtype=torch.int32, device="cuda")
        n = simd.detect_peaks_device(y, out, simd.EXTREMUM_MAXIMUM)
        torch.cuda.synchronize()
    yh = y.cpu().numpy()
    want_p, want_v = oracle.detect_peaks(yh, 1)
    assert n == want_p.size
    assert np.array_equal(out[:n, 0].cpu().numpy(), want_p)


def test_minmax1D_parity(simd, oracle):
    """minmax1D through the C-ABI: bit-exact (min / max are exact) for host and device input,
    lengths around the vector / tile boundaries, NULL result pointers."""
    import ctypes as C
    import torch
    simd.use_torch_stream()
    a = np.ones(100, np.float32)
    a[[0, 1, 10, 89, 21]] = [127, 15, 252, 31, 3]          # tests/normalize.cc:65-81
    assert simd.minmax1D(a) == (1.0, 252.0)
    mx = C.c_float()
    simd.lib.minmax1D(1, a.ctypes.data, 100, None, C.byref(mx))
    assert mx.value == 252.0
    simd.lib.minmax1D(1, a.ctypes.data, 100, None, None)
    rng = np.random.default_rng(6)
    for n in (1, 2, 3, 5, 1023, 1024, 1025, 300001, 5_000_003):
        x = rng.standard_normal(n).astype(np.float32)
        want = oracle.minmax1D(x)
        assert simd.minmax1D(x) == want
        assert simd.minmax1D(x[1:]) == oracle.minmax1D(x[1:]) if n > 1 else True   # misaligned
        xd = torch.from_numpy(x).cuda()
        assert simd.minmax1D(xd) == want
        out = torch.zeros(2, device="cuda")
        simd.lib.simd_b200_minmax1D_device(xd.data_ptr(), n, out.data_ptr())
        torch.cuda.synchronize()
        assert tuple(out.cpu().tolist()) == want
    # NaN samples.  The reference's scalar loop (src/normalize.c:425-443) starts from src[0] and
    # updates with `<` / `>`, so NaNs after the first sample never win a comparison: they are
    # ignored -- and so they are here (fminf / fmaxf), for any position but the first.  A NaN in
    # src[0] makes the reference return NaN for both; here it is ignored like any other (pinned
    # below: the one documented difference, include/simd/normalize.h).
    x = rng.standard_normal(100001).astype(np.float32)
    x[[7, 5000, 99999]] = np.nan
    want = oracle.minmax1D(x)
    assert not np.isnan(want).any()
    assert simd.minmax1D(x) == want
    assert simd.minmax1D(torch.from_numpy(x).cuda()) == want
    x[0] = np.nan
    assert simd.minmax1D(x) == (float(np.nanmin(x)), float(np.nanmax(x)))


def test_int16_converters_parity(simd, oracle):
    """int16_to_float / float_to_int16 through the C-ABI: bit-exact, host and device buffers,
    aligned and misaligned starts, saturation."""
    import torch
    simd.use_torch_stream()
    rng = np.random.default_rng(12)
    for n in (1, 7, 8, 9, 4099, 1_000_003):
        for off in (0, 1):
            i16 = rng.integers(-32768, 32768, n + off, dtype=np.int16)[off:]
            out = np.empty(n, np.float32)
            simd.lib.int16_to_float(i16.ctypes.data, n, out.ctypes.data)
            assert np.array_equal(out, oracle.int16_to_float(i16))
            x = rng.uniform(-40000, 40000, n + off).astype(np.float32)[off:]
            q = np.empty(n, np.int16)
            simd.lib.float_to_int16(x.ctypes.data, n, q.ctypes.data)
            assert np.array_equal(q, oracle.float_to_int16(x))
    i16 = torch.from_numpy(rng.integers(-32768, 32768, 100000, dtype=np.int16)).cuda()
    f = torch.empty(100000, device="cuda")
    simd.lib.int16_to_float(i16.data_ptr(), 100000, f.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(f, i16.float())
    # the chain the converters exist for: int16 PCM in, correlate on device
    h = uni(5, 200)
    with simd.Handle(100000, 200, "", True) as hd:
        y = hd(f, torch.from_numpy(h).cuda())
        torch.cuda.synchronize()
    want = oracle.convolve_fft(f.cpu().numpy(), h, 1)
    check(y.cpu().numpy(), want, f.cpu().numpy(), h, "int16 -> correlate")


def test_kernel_launch_counter(simd):
    before = simd.lib.simd_b200_kernel_launches()
    simd.convolve_simd(KAT_X, KAT_H)
    assert simd.lib.simd_b200_kernel_launches() == before + 1
