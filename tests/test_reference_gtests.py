"""The reference's own gtests (/root/reference/tests/convolve.cc:53-166 and
tests/correlate.cc:53-152, with its bundled gtest) compiled against include/ and linked to
the product library -- the direct proof that a C/C++ translation unit written for the
reference runs unchanged on libSimd.so (B200).

The binaries are built by tests/ref_gtests/Makefile from the sources where they lie (run by
__graft_entry__.build() when /root/reference is present); they travel to the GPU box with
the snapshot.  8 tests: convolve_reference (known answer), convolve_fft,
convolve_overlap_save, convolve_simd and the four cross_correlate twins; the checks are the
reference's own (ASSERT_NEAR 1e-4 on the known answers, squared difference <= 1e-6 /
ASSERT_NEAR 1e-3 against its simd=0 call).
"""
import os
import re
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "ref_gtests", "_bin")

EXPECTED = {
    "ref_gtest_convolve": ["convolve.convolve_reference", "convolve.convolve_fft",
                           "convolve.convolve_overlap_save", "convolve.convolve_simd"],
    "ref_gtest_correlate": ["correlate.cross_correlate_reference", "correlate.cross_correlate_fft",
                            "correlate.cross_correlate_overlap_save", "correlate.cross_correlate_simd"],
}


def _binary(name):
    path = os.path.join(BIN, name)
    if not os.path.exists(path):
        pytest.skip(f"{path} not built (needs /root/reference; __graft_entry__.build() makes it)")
    return path


def test_binaries_link_the_product_library_only():
    """CPU-side: the test binaries resolve every convolve/correlate symbol from libSimd.so (no
    oracle, no reference object linked in)."""
    for name in EXPECTED:
        path = _binary(name)
        needed = subprocess.run(["readelf", "-d", path], capture_output=True, text=True, check=True).stdout
        assert "libSimd.so" in needed
        assert "oracle" not in needed and "libSimd_ref" not in needed
        syms = subprocess.run(["nm", "-D", "--undefined-only", path], capture_output=True, text=True,
                              check=True).stdout
        prefix = "convolve" if name.endswith("convolve") else "cross_correlate"
        for s in (prefix + "_simd", prefix + "_fft_initialize", prefix + "_overlap_save"):
            assert re.search(rf"\bU {s}\b", syms), (name, s)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(EXPECTED))
def test_reference_gtests_pass_on_libsimd(name):
    path = _binary(name)
    r = subprocess.run([path], capture_output=True, text=True, timeout=120)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-3000:]
    for t in EXPECTED[name]:
        assert re.search(rf"\[\s+OK \] {re.escape(t)} ", out), (t, out[-3000:])
    assert "[  PASSED  ] 4 tests." in out, out[-1500:]
    assert "FAILED" not in out
