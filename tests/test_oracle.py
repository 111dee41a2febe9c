"""CPU tests: pin the oracle (C restatement) against the reference's known answers, the
committed golden fixtures (generated from the reference's own sources, oracle/_ref) and,
when it is built here, oracle/_ref itself."""
import numpy as np
import pytest

from _oracle import parity_ok, rel_l2

KAT_CONV = [10, 29, 56, 90, 124, 158, 192, 226, 170, 113, 56]     # tests/convolve.cc:53-71
KAT_CORR = [7, 22, 46, 80, 114, 148, 182, 216, 187, 142, 80]      # tests/correlate.cc:53-71


def test_known_answers(oracle):
    x = np.arange(1, 9, dtype=np.float32)
    h = np.array([10, 9, 8, 7], dtype=np.float32)
    assert np.allclose(oracle.convolve_brute(x, h), KAT_CONV, atol=1e-4)
    assert np.allclose(oracle.cross_correlate_brute(x, h), KAT_CORR, atol=1e-4)
    assert np.allclose(oracle.convolve_f64(x, h), KAT_CONV)
    assert np.allclose(oracle.convolve_f64(x, h, 1), KAT_CORR)


def test_correlate_is_convolve_with_reversed_h(oracle):
    rng = np.random.default_rng(3)
    x, h = rng.uniform(-1, 1, 300).astype(np.float32), rng.uniform(-1, 1, 17).astype(np.float32)
    a = oracle.cross_correlate_brute(x, h)
    b = oracle.convolve_brute(x, h[::-1].copy())
    assert rel_l2(a, b) < 1e-6


def test_golden_brute_force_bit_exact(oracle, golden):
    """The restatement's scalar loops reproduce the reference's simd=0 path bit for bit."""
    n = 0
    for name, c in golden.items():
        if c["path"] != "simd0":
            continue
        fn = oracle.cross_correlate_brute if c["op"] == "correlate" else oracle.convolve_brute
        assert np.array_equal(fn(c["x"], c["h"]), c["y"]), name
        n += 1
    assert n >= 20


def test_golden_all_paths_within_tolerance(oracle, golden):
    """Every reference path (AVX, FFT, overlap-save, selector) vs the oracle's matching
    path and vs the float64 anchor, at the north-star tolerance."""
    for name, c in golden.items():
        x, h, y = c["x"], c["h"], c["y"]
        rev = c["op"] == "correlate"
        anchor = oracle.convolve_f64(x, h, rev)
        ok, r, m = parity_ok(y, anchor, x, h)
        assert ok, (name, r, m)
        if c["path"] == "fft":
            mine = oracle.convolve_fft(x, h, rev)
        elif c["path"] == "overlap_save":
            mine = oracle.convolve_overlap_save(x, h, rev)
        elif c["path"] == "auto":
            mine = oracle.cross_correlate(x, h) if rev else oracle.convolve(x, h)
            assert oracle.select_algorithm(x.size, h.size) == c["algorithm"], name
            if c["algorithm"] == 2:
                assert oracle.overlap_save_length(h.size) == c["transform"], name
            elif c["algorithm"] == 1:
                assert oracle.fft_length(x.size, h.size) == c["transform"], name
        else:
            continue
        ok, r, m = parity_ok(mine, y, x, h, rel=2e-6)
        assert ok, (name, r, m)


def test_reference_tests_shapes(oracle, golden):
    """tests/convolve.cc:73-166, tests/correlate.cc:73-152: delta^2 <= 1e-6 / +-1e-3."""
    for xl, path in ((1020, "fft"), (1021, "os"), (1024, "simd1")):
        for op in ("convolve", "correlate"):
            c = golden[f"sin{xl}_{op}_{path}"]
            v = golden[f"sin{xl}_{op}_simd0"]["y"]
            assert np.max((c["y"] - v) ** 2) <= 1e-6
            rev = op == "correlate"
            mine = {"fft": oracle.convolve_fft, "os": oracle.convolve_overlap_save}.get(path)
            if mine is not None:
                assert np.max((mine(c["x"], c["h"], rev) - v) ** 2) <= 1e-6


def test_selector_boundaries(oracle):
    # SURVEY 3.1 [probed]: reference selector decisions
    assert oracle.select_algorithm(200, 50) == 0 and oracle.select_algorithm(201, 50) == 2
    assert oracle.select_algorithm(350, 350) == 0 and oracle.select_algorithm(351, 350) == 1
    assert oracle.select_algorithm(1024, 50) == 2 and oracle.overlap_save_length(50) == 128
    assert oracle.overlap_save_length(512) == 2048 and oracle.overlap_save_length(64) == 256
    assert oracle.overlap_save_length(4096) == 16384 and oracle.overlap_save_length(8192) == 32768
    assert oracle.fft_length(351, 350) == 1024 and oracle.fft_length(512, 512) == 1024


def test_oracle_against_reference_build(oracle, reference):
    """Live comparison with the reference's own code (only where oracle/_ref is built)."""
    rng = np.random.default_rng(11)
    for xl, hl in [(1, 1), (5, 3), (64, 64), (1000, 37), (4097, 129), (20000, 1000)]:
        x = rng.uniform(-1, 1, xl).astype(np.float32)
        h = rng.uniform(-1, 1, hl).astype(np.float32)
        assert np.array_equal(oracle.convolve_brute(x, h), reference.simd(x, h, simd=0))
        assert np.array_equal(oracle.cross_correlate_brute(x, h),
                              reference.simd(x, h, correlate=True, simd=0))
        assert rel_l2(reference.simd(x, h, simd=1), oracle.convolve_brute(x, h)) < 2e-6
        if hl < xl // 2:
            xg = np.zeros(xl + 4096, np.float32)
            xg[:xl] = x
            a = reference.handle_call("_overlap_save", xg[:xl], h)
            assert rel_l2(oracle.convolve_overlap_save(x, h), a) < 2e-6
        if xl + hl - 1 >= 8:   # the reference's AVX multiply overruns tiny buffers
            a = reference.handle_call("_fft", x, h, correlate=True)
            assert rel_l2(oracle.convolve_fft(x, h, 1), a) < 2e-6


def test_reference_fft_handle_is_single_use(reference, oracle):
    """documents why parity runs use a fresh reference handle per call."""
    rng = np.random.default_rng(5)
    x, h = rng.uniform(-1, 1, 600).astype(np.float32), rng.uniform(-1, 1, 400).astype(np.float32)
    outs = reference.handle_call("_fft", x, h, calls=2)
    want = oracle.convolve_f64(x, h)
    assert rel_l2(outs[0], want) < 1e-5
    assert rel_l2(outs[1], want) > 1e-3


def test_oracle_batch_threads(oracle):
    rng = np.random.default_rng(9)
    x = rng.uniform(-1, 1, (6, 3000)).astype(np.float32)
    h = rng.uniform(-1, 1, 100).astype(np.float32)
    out, used = oracle.convolve_batch(x, h, threads=2)
    assert used == 2
    for b in range(6):
        assert rel_l2(out[b], oracle.convolve_f64(x[b], h)) < 1e-5


def test_detect_peaks_oracle_against_golden_and_reference(oracle, reference_or_none):
    """SURVEY 8(f) rank 1.  tests/detect_peaks.cc:43-100 known answers + fixtures from the
    reference's scalar loop; bit-exact (index work)."""
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "peaks.npz"))
    names = sorted({k.split("/")[0] for k in z.files})
    assert {"sin", "nasty", "noise", "corr_c2"} <= set(names)
    for name in names:
        data = z[name + "/data"]
        for t in (1, 2, 3):
            pos, val = oracle.detect_peaks(data, t)
            assert np.array_equal(pos, z[f"{name}/pos{t}"]), (name, t)
            assert np.array_equal(val, z[f"{name}/val{t}"]), (name, t)
            if reference_or_none is not None:
                rp, rv = reference_or_none.detect_peaks(data, t, simd=0)
                assert np.array_equal(pos, rp) and np.array_equal(val, rv)
    assert z["sin/pos1"].tolist() == [i * 200 + 50 for i in range(20)]
    assert z["sin/pos2"].tolist() == [i * 200 + 150 for i in range(20)]
    assert z["nasty/pos1"].tolist() == [7, 16, 97, 99]


def test_minmax1D_oracle(oracle, reference_or_none):
    """SURVEY 8(f) rank 3.  tests/normalize.cc:65-81 known answer + the reference build."""
    a = np.ones(100, np.float32)
    a[[0, 1, 10, 89, 21]] = [127, 15, 252, 31, 3]
    assert oracle.minmax1D(a) == (1.0, 252.0)
    rng = np.random.default_rng(4)
    for n in (1, 7, 8, 9, 1000, 100003):
        x = rng.standard_normal(n).astype(np.float32)
        assert oracle.minmax1D(x) == (float(x.min()), float(x.max()))
        if reference_or_none is not None:
            assert reference_or_none.minmax1D(x, 1) == oracle.minmax1D(x)
            assert reference_or_none.minmax1D(x, 0) == oracle.minmax1D(x)


def test_int16_converters_oracle(oracle, reference_or_none):
    """plain casts (inc/simd/arithmetic.h:43-57), bit-exact."""
    rng = np.random.default_rng(2)
    i16 = rng.integers(-32768, 32768, 10007, dtype=np.int16)
    f = oracle.int16_to_float(i16)
    assert np.array_equal(f, i16.astype(np.float32))
    x = (rng.uniform(-32768, 32767, 10007)).astype(np.float32)
    x[:6] = [0.9, -0.9, 1.5, -1.5, 32767.9 - 1, -32768.0]
    q = oracle.float_to_int16(x)
    assert np.array_equal(q, np.trunc(x).astype(np.int16))
    assert oracle.float_to_int16(np.array([1e9, -1e9, np.nan], np.float32)).tolist() == [32767, -32768, 0]
    if reference_or_none is not None:
        assert np.array_equal(reference_or_none.int16_to_float(i16), f)
        assert np.array_equal(reference_or_none.float_to_int16(x), q)
