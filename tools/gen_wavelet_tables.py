#!/usr/bin/env python
"""tools/gen_wavelet_tables.py -- computes the wavelet filter tables from first principles and
writes veles/simd_b200/src/wavelet_tables.c.

The reference ships three literature tables (/root/reference/src/daubechies.c, symlets.c,
coiflets.c: lowpass filters for Daubechies 2..76 taps, Symlets 2..76 taps, Coiflets 6..30
taps).  Nothing is taken from those files; the filters are recomputed here with mpmath:

  Daubechies p (2p taps)  spectral factorisation of the Daubechies polynomial
                          P(y) = sum_k C(p-1+k, k) y^k,  y = (2 - z - 1/z)/4, keeping of every
                          reciprocal root pair {z, 1/z} the root outside the unit circle;
                          normalised to sum sqrt(2) (the reference's convention for this family).
  Symlet p                the same polynomial, but a mix of inside/outside roots ("least
                          asymmetric" choice).  The choice is a discrete parameter: one bit per
                          root group, SYMLET_SELECTION below, in the canonical group order
                          (groups sorted by Re(y), Im(y) of the y-root with Im >= 0; 1 = the
                          root inside the unit circle).  `--derive` recovers the bits by testing
                          which member of each pair is a zero of the reference's filter (needs
                          oracle/_ref/libSimd_ref.so, i.e. this container); the filters
                          themselves are then recomputed exactly.  Normalised to sum 1.
  Coiflet K (6K taps)     Gauss-Newton on the defining equations (orthonormality, 2K vanishing
                          wavelet moments, 2K-1 vanishing shifted scaling moments about n = 2K)
                          from a two-significant-digit starting guess; normalised to sum 1.

The reference's float tables are its double tables printed with 8 significant digits
(`%.7e`) and parsed as float; the same rounding is applied here, so the Daubechies and Symlet
rows are bit-identical to the reference's kDaubechiesF / kSymletsF (tests/test_wavelet_tables.py
checks that against oracle/_ref when it is present, and against committed digests otherwise).
The reference's Coiflet rows for 24 and 30 taps carry the 12-digit literature values, which
are off the exact filters by 1.7e-8 and 7.8e-6; this table holds the exact filters.
"""
import argparse
import os
import sys
from fractions import Fraction

import mpmath as mp
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
mp.mp.dps = 120
EPS = mp.mpf(10) ** -60

# one string per Symlet p = 1..38; see the module docstring
SYMLET_SELECTION = [
    "", "0", "0", "01", "10", "101", "100", "0101", "0110", "10101", "01100", "101010", "001110",
    "0011010", "0011100", "10011010", "01110001", "101100101", "001011100", "1010011010",
    "0011110001", "00110011010", "00110011100", "100111000110", "010011110001", "1011000111001",
    "0101110100011", "10100111000110", "00101101110001", "000111011000110", "000111111000011",
    "0001011111000101", "0010011110110001", "10010011111000101", "00011101111000011",
    "001001101111000101", "010111100001001110", "1001101101000111001",
]


def y_roots(p):
    """Roots of P(y) = sum_k C(p-1+k, k) y^k with Im >= 0, in canonical order."""
    if p == 1:
        return []
    coeffs = [mp.binomial(p - 1 + k, k) for k in range(p)]
    ys = mp.polyroots(coeffs[::-1], maxsteps=4000, extraprec=4000)
    ys = [y for y in ys if mp.im(y) > -EPS]
    ys.sort(key=lambda y: (mp.re(y), mp.im(y)))
    return ys


def z_pair(y):
    """The reciprocal roots of z^2 - (2 - 4y) z + 1: (inside the unit circle, outside)."""
    b = 2 - 4 * y
    d = mp.sqrt(b * b - 4)
    z1, z2 = (b + d) / 2, (b - d) / 2
    return (z1, z2) if abs(z1) < abs(z2) else (z2, z1)


def build_filter(p, ys, selection):
    """((1 + z)/2)^p * prod over groups of the chosen root factor(s), each normalised to 1
    at z = 1; coefficient i multiplies z^i.  Returns mpf coefficients summing to 1."""
    poly = [mp.mpf(1)]
    for _ in range(p):
        poly = [(a + b) / 2 for a, b in zip(poly + [0], [0] + poly)]
    for y, bit in zip(ys, selection):
        zi, zo = z_pair(y)
        r = zi if bit == "1" else zo
        if abs(mp.im(y)) > EPS:        # complex group: (z - r)(z - conj r), real quadratic
            f = [abs(r) ** 2, -2 * mp.re(r), mp.mpf(1)]
        else:
            f = [-mp.re(r), mp.mpf(1)]
        s = sum(f)
        f = [c / s for c in f]
        new = [mp.mpf(0)] * (len(poly) + len(f) - 1)
        for i, a in enumerate(poly):
            for j, c in enumerate(f):
                new[i + j] += a * c
        poly = new
    return poly


def coiflet(K, seed):
    """Gauss-Newton on the Coiflet equations; moments are scaled by the filter length to keep
    the Jacobian well conditioned."""
    n, c = 6 * K, 2 * K
    h = [mp.mpf(s) for s in seed]
    for _ in range(60):
        r, J = [sum(h) - 1], [[mp.mpf(1)] * n]
        for m in range(3 * K):
            r.append(sum(h[i] * h[i + 2 * m] for i in range(n - 2 * m)) - (mp.mpf(1) / 2 if m == 0 else 0))
            row = [mp.mpf(0)] * n
            for i in range(n - 2 * m):
                row[i] += h[i + 2 * m]
                row[i + 2 * m] += h[i]
            J.append(row)
        for j in range(2 * K):
            w = [((-1) ** i) * (mp.mpf(i - c) / n) ** j for i in range(n)]
            r.append(sum(w[i] * h[i] for i in range(n)))
            J.append(w)
        for j in range(1, 2 * K):
            w = [(mp.mpf(i - c) / n) ** j for i in range(n)]
            r.append(sum(w[i] * h[i] for i in range(n)))
            J.append(w)
        d = mp.qr_solve(mp.matrix(J), -mp.matrix(r))[0]
        h = [h[i] + d[i] for i in range(n)]
        if max(abs(d[i]) for i in range(n)) < mp.mpf(10) ** -70:
            return h
    raise RuntimeError("Coiflet %d did not converge" % K)


# two significant digits: starting guesses only, the solve above does the rest
COIFLET_SEED = {
    1: [-5.1e-02, 2.4e-01, 6.0e-01, 2.7e-01, -5.1e-02, -1.1e-02],
    2: [1.2e-02, -2.9e-02, -4.8e-02, 2.7e-01, 5.7e-01, 2.9e-01, -5.4e-02, -4.2e-02, 1.7e-02, 4.0e-03,
        -1.3e-03, -5.1e-04],
    3: [-2.7e-03, 5.5e-03, 1.7e-02, -4.7e-02, -4.3e-02, 2.9e-01, 5.6e-01, 3.0e-01, -5.1e-02, -5.8e-02,
        2.4e-02, 1.1e-02, -6.4e-03, -1.8e-03, 7.9e-04, 3.3e-04, -5.0e-05, -2.4e-05],
    4: [6.3e-04, -1.2e-03, -5.2e-03, 1.1e-02, 1.9e-02, -5.7e-02, -4.0e-02, 2.9e-01, 5.5e-01, 3.1e-01,
        -4.7e-02, -6.8e-02, 2.8e-02, 1.8e-02, -1.1e-02, -4.0e-03, 2.7e-03, 9.0e-04, -4.2e-04, -1.8e-04,
        4.4e-05, 2.2e-05, -2.3e-06, -1.3e-06],
    5: [-1.5e-04, 2.5e-04, 1.5e-03, -2.9e-03, -7.2e-03, 1.7e-02, 2.0e-02, -6.5e-02, -3.7e-02, 3.0e-01,
        5.5e-01, 3.1e-01, -4.4e-02, -7.5e-02, 2.9e-02, 2.3e-02, -1.4e-02, -6.5e-03, 4.8e-03, 1.7e-03,
        -1.2e-03, -4.5e-04, 2.1e-04, 9.9e-05, -2.9e-05, -1.5e-05, 2.6e-06, 1.5e-06, -1.2e-07, -6.7e-08],
}


def float32_like_reference(v):
    """The reference's float rows are its double rows written with 8 significant digits and
    parsed as float: round the decimal string to the nearest float32 exactly."""
    q = Fraction("%.7e" % float(v))
    c = np.float32(float(q))
    best = c
    for cand in (np.nextafter(c, np.float32(-np.inf)), np.nextafter(c, np.float32(np.inf))):
        if abs(Fraction(float(cand)) - q) < abs(Fraction(float(best)) - q):
            best = cand
    return best


# ---- derivation of the discrete parameters from the reference library (this container) ----
def _ref_tables():
    import ctypes
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libSimd_ref.so"))
    sym = np.ctypeslib.as_array((ctypes.c_double * 76 * 38).in_dll(lib, "kSymletsD"))
    coif = np.ctypeslib.as_array((ctypes.c_double * 30 * 5).in_dll(lib, "kCoifletsD"))
    return sym, coif


def derive():
    """Prints SYMLET_SELECTION and COIFLET_SEED.  For every root group the member that is a
    zero of the reference's Symlet filter is found by evaluating that filter at both
    candidates (relative to the evaluation's rounding scale); groups where the 16-digit table
    cannot tell are settled by trying both and comparing the rebuilt filter."""
    import itertools
    sym, coif = _ref_tables()

    def pv(c, z):
        s = mp.mpc(0)
        for a in reversed(c):
            s = s * z + a
        return abs(s)

    def scale(c, z):
        s, a = mp.mpf(0), abs(z)
        for v in reversed(c):
            s = s * a + abs(v)
        return s

    def factor64(r, cplx):
        r = complex(r)
        f = np.array([abs(r) ** 2, -2 * r.real, 1.0]) if cplx else np.array([-r.real, 1.0])
        return f / f.sum()

    print("SYMLET_SELECTION = [")
    for p in range(1, 39):
        want = sym[p - 1][:2 * p]
        ref = [mp.mpf(float(v)) for v in want]
        ys = y_roots(p)
        fixed, amb = {}, []
        for k, y in enumerate(ys):
            zi, zo = z_pair(y)
            si = mp.log10(pv(ref, zi) / scale(ref, zi) + mp.mpf(10) ** -100)
            so = mp.log10(pv(ref, zo) / scale(ref, zo) + mp.mpf(10) ** -100)
            if abs(si - so) > 5:
                fixed[k] = "1" if si < so else "0"
            else:
                amb.append(k)
        base = np.array([1.0])
        for _ in range(p):
            base = np.convolve(base, [0.5, 0.5])
        for k, bit in fixed.items():
            zi, zo = z_pair(ys[k])
            base = np.convolve(base, factor64(zi if bit == "1" else zo, abs(mp.im(ys[k])) > EPS))
        best = None
        for combo in itertools.product("10", repeat=len(amb)):
            h = base
            for k, bit in zip(amb, combo):
                zi, zo = z_pair(ys[k])
                h = np.convolve(h, factor64(zi if bit == "1" else zo, abs(mp.im(ys[k])) > EPS))
            e = np.abs(h - want).max()
            if best is None or e < best[0]:
                best = (e, combo)
        sel = dict(fixed)
        sel.update(dict(zip(amb, best[1])))
        s = "".join(sel[k] for k in range(len(ys)))
        exact = np.array([float(c) for c in build_filter(p, ys, s)])
        print('    "%s",  # %d: max |exact - reference double row| = %.1e' % (s, p, np.abs(exact - want).max()),
              flush=True)
    print("]")
    print("COIFLET_SEED = {")
    for K in range(1, 6):
        print("    %d: [%s]," % (K, ", ".join("%.1e" % v for v in coif[K - 1][:6 * K])))
    print("}")


# ---- output ---------------------------------------------------------------------------------
def compute_tables():
    daub, syml, coif = [], [], []
    for p in range(1, 39):
        ys = y_roots(p)
        d = build_filter(p, ys, "0" * len(ys))
        daub.append([c * mp.sqrt(2) for c in d])
        syml.append(build_filter(p, ys, SYMLET_SELECTION[p - 1]))
        print("order %d done" % (2 * p), file=sys.stderr, flush=True)
    for K in range(1, 6):
        coif.append(coiflet(K, COIFLET_SEED[K]))
    return daub, syml, coif


def emit_rows(f, name, rows, width):
    f.write("const float %s[%d][%d] = {\n" % (name, len(rows), width))
    for k, row in enumerate(rows):
        f.write("  /* %d taps */ {\n" % len(row))
        vals = [float32_like_reference(v) for v in row]
        for i in range(0, len(vals), 4):
            f.write("    " + " ".join("%sf," % float(v).hex() for v in vals[i:i + 4]) + "\n")
        f.write("  },\n")
    f.write("};\n\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--derive", action="store_true",
                    help="print the discrete parameters recovered from oracle/_ref (this container only)")
    ap.add_argument("-o", default=os.path.join(ROOT, "veles", "simd_b200", "src", "wavelet_tables.c"))
    args = ap.parse_args()
    if args.derive:
        derive()
        return
    daub, syml, coif = compute_tables()
    with open(args.o, "w") as f:
        f.write("/* wavelet_tables.c -- GENERATED by tools/gen_wavelet_tables.py; do not edit.\n"
                " * Lowpass filters computed from first principles (see the generator's docstring):\n"
                " * Daubechies 2..76 taps (sum sqrt 2), Symlets 2..76 taps (sum 1), Coiflets 6..30 taps\n"
                " * (sum 1), rounded to float the way the reference rounds its own tables\n"
                " * (/root/reference/src/daubechies.c:1595, symlets.c, coiflets.c: 8 significant digits).\n"
                " * Hexadecimal float literals: exact, no decimal round trip. */\n"
                "#include \"wavelet_tables.h\"\n\n")
        emit_rows(f, "b200_daubechies", daub, 76)
        emit_rows(f, "b200_symlets", syml, 76)
        emit_rows(f, "b200_coiflets", coif, 30)
    print("wrote", args.o)


if __name__ == "__main__":
    main()
