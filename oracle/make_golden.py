#!/usr/bin/env python
"""Generate tests/golden/golden.npz from the reference's OWN code (oracle/_ref, i.e.
/root/reference/src/{convolve,correlate,memory}.c compiled against the FFT shim).

TEST INFRASTRUCTURE.  Run in the build container (needs /root/reference to have built
oracle/_ref/libSimd_ref.so):   python oracle/make_golden.py
The fixtures are committed; the GPU box and the CPU test-suite only read them.

Each case stores x, h, the op ("convolve" | "correlate"), the reference entry point that
produced `y` ("simd0" scalar loop, "simd1" AVX loop, "fft", "overlap_save", "auto" = the
convolve_initialize / cross_correlate_initialize selector path) and the handle facts the
reference reported (algorithm, L or M).
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tests"))
from _oracle import Reference  # noqa: E402

ref = Reference()
cases = {}


def add(name, x, h, op, path, guard=0):
    x = np.ascontiguousarray(x, dtype=np.float32)
    h = np.ascontiguousarray(h, dtype=np.float32)
    corr = op == "correlate"
    xin = x
    if guard:  # reference over-reads x in the first overlap-save block
        buf = np.zeros(x.size + guard, dtype=np.float32)
        buf[:x.size] = x
        xin = buf[:x.size]
    info = {}
    if path == "simd0":
        y = ref.simd(xin, h, correlate=corr, simd=0)
    elif path == "simd1":
        y = ref.simd(xin, h, correlate=corr, simd=1)
    else:
        kind = {"fft": "_fft", "overlap_save": "_overlap_save", "auto": ""}[path]
        y, info = ref.handle_call(kind, xin, h, correlate=corr, return_handle_info=True)
    cases[name] = dict(x=x, h=h, y=y, op=op, path=path, info=info)


# --- known-answer tests: tests/convolve.cc:53-71, tests/correlate.cc:53-71
kx = np.arange(1, 9, dtype=np.float32)
kh = np.array([10, 9, 8, 7], dtype=np.float32)
add("kat_convolve", kx, kh, "convolve", "simd0")
add("kat_correlate", kx, kh, "correlate", "simd0")
assert cases["kat_convolve"]["y"].tolist() == [10, 29, 56, 90, 124, 158, 192, 226, 170, 113, 56]
assert cases["kat_correlate"]["y"].tolist() == [7, 22, 46, 80, 114, 148, 182, 216, 187, 142, 80]

# --- the reference tests' synthetic pattern: tests/convolve.cc:74-84,107-117,140-150
def ref_pattern(xlen, hlen=50):
    i = np.arange(xlen, dtype=np.float32)
    x = (np.sin(i).astype(np.float32) * np.float32(100)).astype(np.float32)
    h = (np.arange(hlen, dtype=np.float32) / np.float32(hlen - 1.0)).astype(np.float32)
    return x, h

for op in ("convolve", "correlate"):
    x, h = ref_pattern(1020); add(f"sin1020_{op}_fft", x, h, op, "fft")
    add(f"sin1020_{op}_simd0", x, h, op, "simd0")
    x, h = ref_pattern(1021); add(f"sin1021_{op}_os", x, h, op, "overlap_save")
    add(f"sin1021_{op}_simd0", x, h, op, "simd0")
    x, h = ref_pattern(1024); add(f"sin1024_{op}_simd1", x, h, op, "simd1")
    add(f"sin1024_{op}_simd0", x, h, op, "simd0")

# --- BASELINE configs that fit a fixture (U(-1,1), seed 0xC0FFEE + config index)
def uni(seed, n):
    return np.random.default_rng(seed).uniform(-1, 1, n).astype(np.float32)

x, h = uni(0xC0FFEE + 1, 1024), uni(0xC1FFEE + 1, 50)
add("c1_convolve_auto", x, h, "convolve", "auto")
add("c1_convolve_simd1", x, h, "convolve", "simd1")
add("c1_convolve_simd0", x, h, "convolve", "simd0")
x, h = uni(0xC0FFEE + 2, 65536), uni(0xC1FFEE + 2, 512)
add("c2_correlate_auto", x, h, "correlate", "auto")
add("c2_correlate_simd0", x, h, "correlate", "simd0")
x, h = uni(0xC0FFEE + 3, 16384), uni(0xC1FFEE + 3, 64)
add("c3_signal_convolve_auto", x, h, "convolve", "auto")
add("c3_signal_convolve_simd0", x, h, "convolve", "simd0")

# --- edge shapes and selector boundaries (src/convolve.c:349-363)
for (xl, hl) in [(1, 1), (8, 1), (7, 7), (64, 64), (200, 50), (201, 50), (350, 350),
                 (351, 350), (512, 512), (37, 5), (1000, 3), (4099, 129)]:
    x, h = uni(1000 + xl, xl), uni(2000 + hl, hl)
    for op in ("convolve", "correlate"):
        add(f"edge_{xl}x{hl}_{op}_auto", x, h, op, "auto", guard=4096)
        add(f"edge_{xl}x{hl}_{op}_simd0", x, h, op, "simd0")
# Q1 shape: selector sends (300,128) to overlap-save, first block over-reads x.
x, h = uni(77, 300), uni(78, 128)
add("q1_300x128_convolve_auto", x, h, "convolve", "auto", guard=4096)
add("q1_300x128_convolve_simd0", x, h, "convolve", "simd0")

# --- detect_peaks: reference scalar path, tests/detect_peaks.cc:43-100
peaks = {}
sin_series = np.sin((np.arange(4000) * np.pi / 100).astype(np.float32).astype(np.float64)).astype(np.float32)
nasty = np.zeros(101, dtype=np.float32)
nasty[[7, 16, 97, 99]] = 1
noise = uni(4242, 100000)
corr = cases["c2_correlate_simd0"]["y"]
for pname, series in (("sin", sin_series), ("nasty", nasty), ("noise", noise), ("corr_c2", corr),
                      ("tiny3", np.array([0, 1, 0], np.float32)), ("flat", np.ones(50, np.float32))):
    peaks[pname + "/data"] = series
    for t in (1, 2, 3):
        pos, val = ref.detect_peaks(series, t, simd=0)
        peaks[f"{pname}/pos{t}"] = pos
        peaks[f"{pname}/val{t}"] = val
assert peaks["sin/pos1"].tolist() == [i * 200 + 50 for i in range(20)]
assert peaks["nasty/pos1"].tolist() == [7, 16, 97, 99]
pdst = os.path.join(os.path.dirname(HERE), "tests", "golden", "peaks.npz")
np.savez_compressed(pdst, **peaks)
print(f"wrote {pdst}: {os.path.getsize(pdst)/1e6:.2f} MB")

out = {}
index = []
for name, c in cases.items():
    out[name + "/x"] = c["x"]; out[name + "/h"] = c["h"]; out[name + "/y"] = c["y"]
    info = c["info"] or {}
    index.append("|".join([name, c["op"], c["path"], str(info.get("algorithm", -1)),
                           str(info.get("L", info.get("M", 0)))]))
out["index"] = np.array(index)
dst = os.path.join(os.path.dirname(HERE), "tests", "golden", "golden.npz")
np.savez_compressed(dst, **out)
print(f"wrote {dst}: {len(cases)} cases, {os.path.getsize(dst)/1e6:.2f} MB")

# ---------------------------------------------------------------------------------------------
# wavelet fixtures: tests/golden/wavelet.npz, produced by the
# reference's own wavelet.c (scalar *_na entry points; the AVX entry point where noted).
# Case name: w|<type>|<order>|<ext>|<level or 0>|<input>, arrays <name>/hi and <name>/lo; the
# inputs are stored once under in/<input>.
import hashlib  # noqa: E402
from _oracle import WAVELET_TYPES, EXTENSIONS  # noqa: E402

wav = {}
rng = np.random.default_rng(0xDA7)
inputs = {
    "ramp32": np.arange(32, dtype=np.float32),                   # tests/wavelet.cc:84-88
    "ramp512": np.arange(512, dtype=np.float32),                 # tests/wavelet.cc:176-183
    "uni1000": rng.uniform(-1, 1, 1000).astype(np.float32),
    "uni4098": rng.uniform(-1, 1, 4098).astype(np.float32),
}
for k, a in inputs.items():
    wav["in/" + k] = a


def wadd(tname, order, ename, level, iname, na=True):
    hi, lo = ref.wavelet_apply(WAVELET_TYPES[tname], order, EXTENSIONS[ename], inputs[iname],
                               level if level else None, na=na)
    name = "w|%s|%d|%s|%d|%s%s" % (tname, order, ename, level, iname, "" if na else "|simd")
    wav[name + "/hi"], wav[name + "/lo"] = hi, lo
    return hi, lo


# known answers of the reference's tests: tests/wavelet.cc:83-108 and :110-166
hi, lo = wadd("daubechies", 8, "periodic", 0, "ramp32")
assert abs(lo[0] - 1.42184071797210) < 1e-6 and abs(lo[15] - 28.8405938767906) < 1e-5
assert abs(hi[13] + 15.5030002317990) < 1e-5
wadd("daubechies", 8, "periodic", 1, "ramp32")
# the reference tests' grids: tests/wavelet.cc:262-300
for tname, orders, levels in (("daubechies", (2, 4, 6, 8, 12, 16), (1, 2, 3, 4)),
                              ("symlet", (2, 4, 6, 8, 12, 16), (1, 2, 3, 4)),
                              ("coiflet", (6, 12), (1, 2))):
    for order in orders:
        for ename in EXTENSIONS:
            wadd(tname, order, ename, 0, "ramp512")
            for level in levels:
                wadd(tname, order, ename, level, "ramp512")
# beyond the reference's grid: long filters, every family's last rows, odd tile remainders
for tname, order in (("daubechies", 20), ("daubechies", 38), ("daubechies", 76), ("symlet", 20),
                     ("symlet", 46), ("symlet", 64), ("symlet", 74), ("symlet", 76),
                     ("coiflet", 18), ("coiflet", 24), ("coiflet", 30)):
    for ename in EXTENSIONS:
        wadd(tname, order, ename, 0, "uni4098")
    wadd(tname, order, "periodic", 0, "uni1000")
    wadd(tname, order, "mirror", 3, "uni1000")
# the AVX entry point, fed the way the reference's tests feed it (tests/wavelet.cc:184-189,
# 229-232): prepared source, destinations from wavelet_allocate_destination
import ctypes as C  # noqa: E402
_fp = C.POINTER(C.c_float)
L = ref.lib
L.wavelet_prepare_array.argtypes = [C.c_int, _fp, C.c_size_t]
L.wavelet_prepare_array.restype = C.c_void_p
L.wavelet_allocate_destination.argtypes = [C.c_int, C.c_size_t]
L.wavelet_allocate_destination.restype = C.c_void_p
L.wavelet_apply.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
L.wavelet_apply.restype = None
libc = C.CDLL(None)
libc.free.argtypes = [C.c_void_p]
for tname, orders in (("daubechies", (2, 4, 6, 8, 12, 16)), ("symlet", (4, 8, 16)), ("coiflet", (6, 12))):
    for order in orders:
        src = inputs["ramp512"]
        prep = L.wavelet_prepare_array(order, src.ctypes.data_as(_fp), src.size)
        dhi = L.wavelet_allocate_destination(order, src.size)
        dlo = L.wavelet_allocate_destination(order, src.size)
        L.wavelet_apply(WAVELET_TYPES[tname], order, EXTENSIONS["mirror"], prep, src.size, dhi, dlo)
        name = "w|%s|%d|mirror|0|ramp512|simd" % (tname, order)
        wav[name + "/hi"] = np.ctypeslib.as_array(C.cast(dhi, _fp), shape=(src.size // 2,)).copy()
        wav[name + "/lo"] = np.ctypeslib.as_array(C.cast(dlo, _fp), shape=(src.size // 2,)).copy()
        for ptr in (prep, dhi, dlo):
            libc.free(ptr)
# digests of the reference's float table rows: pins the product's generated tables on machines
# without the reference (the rows themselves are not stored)
for tname, tid in WAVELET_TYPES.items():
    table = ref.wavelet_table(tid)
    step = 6 if tname == "coiflet" else 2
    wav["digest/" + tname] = np.array([hashlib.sha256(table[k][:(k + 1) * step].tobytes()).hexdigest()
                                       for k in range(table.shape[0])])
wdst = os.path.join(os.path.dirname(HERE), "tests", "golden", "wavelet.npz")
np.savez_compressed(wdst, **wav)
print(f"wrote {wdst}: {sum(1 for k in wav if k.endswith('/hi'))} cases, {os.path.getsize(wdst)/1e6:.2f} MB")
