"""ctypes access to the CPU oracle (oracle/liboracle.so) and, when present, to the
reference's own sources compiled here (oracle/_ref/libSimd_ref.so).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package.
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "liboracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libSimd_ref.so")

_fp = C.POINTER(C.c_float)
_dp = C.POINTER(C.c_double)


def build():
    """(Re)build liboracle.so and, when /root/reference exists, oracle/_ref."""
    subprocess.run(["make", "-C", ORACLE_DIR, "--no-print-directory", "-s"], check=True)


def _f32(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(_fp)


class Oracle:
    """The C restatement (kind == "port")."""

    def __init__(self):
        if not os.path.exists(ORACLE_SO):
            build()
        self.lib = L = C.CDLL(ORACLE_SO)
        sz = C.c_size_t
        L.oracle_convolve_brute.argtypes = [_fp, sz, _fp, sz, _fp]
        L.oracle_cross_correlate_brute.argtypes = [_fp, sz, _fp, sz, _fp]
        L.oracle_convolve_f64.argtypes = [_fp, sz, _fp, sz, C.c_int, _dp]
        L.oracle_overlap_save_length.argtypes = [sz]
        L.oracle_fft_length.argtypes = [sz, sz]
        L.oracle_select_algorithm.argtypes = [sz, sz]
        L.oracle_convolve_overlap_save.argtypes = [_fp, sz, _fp, sz, C.c_int, _fp]
        L.oracle_convolve_fft.argtypes = [_fp, sz, _fp, sz, C.c_int, _fp]
        L.oracle_convolve.argtypes = [_fp, sz, _fp, sz, _fp]
        L.oracle_cross_correlate.argtypes = [_fp, sz, _fp, sz, _fp]
        L.oracle_convolve_batch.argtypes = [_fp, sz, sz, _fp, sz, _fp, sz, sz, C.c_int, C.c_int]
        L.oracle_convolve_batch.restype = C.c_int
        L.oracle_detect_peaks.argtypes = [_fp, sz, C.c_int, C.POINTER(C.c_int), _fp, sz]
        L.oracle_detect_peaks.restype = sz
        L.oracle_minmax1D.argtypes = [_fp, C.c_int, _fp, _fp]
        L.oracle_minmax1D.restype = None
        L.oracle_int16_to_float.argtypes = [C.c_void_p, sz, _fp]
        L.oracle_float_to_int16.argtypes = [_fp, sz, C.c_void_p]
        L.oracle_wavelet_apply.argtypes = [_fp, C.c_int, C.c_int, _fp, sz, _fp, _fp]
        L.oracle_wavelet_apply.restype = None
        L.oracle_stationary_wavelet_apply.argtypes = [_fp, C.c_int, C.c_int, C.c_int, _fp, sz, _fp, _fp]
        L.oracle_stationary_wavelet_apply.restype = None

    def wavelet_apply(self, row, ext, src, level=None):
        """(desthi, destlo) of the reference's scalar loops for the lowpass table row `row`
        (order = row.size); level None = decimated transform, else stationary at that level."""
        row, rp = _f32(row)
        src, sp = _f32(src)
        n_out = src.size // 2 if level is None else src.size
        hi, lo = np.empty(n_out, np.float32), np.empty(n_out, np.float32)
        if level is None:
            self.lib.oracle_wavelet_apply(rp, row.size, int(ext), sp, src.size,
                                          hi.ctypes.data_as(_fp), lo.ctypes.data_as(_fp))
        else:
            self.lib.oracle_stationary_wavelet_apply(rp, row.size, int(level), int(ext), sp, src.size,
                                                     hi.ctypes.data_as(_fp), lo.ctypes.data_as(_fp))
        return hi, lo

    def _run(self, fn, x, h, *extra):
        x, xp = _f32(x)
        h, hp = _f32(h)
        out = np.empty(x.size + h.size - 1, dtype=np.float32)
        fn(xp, x.size, hp, h.size, *extra, out.ctypes.data_as(_fp))
        return out

    def convolve_brute(self, x, h):
        return self._run(self.lib.oracle_convolve_brute, x, h)

    def cross_correlate_brute(self, x, h):
        return self._run(self.lib.oracle_cross_correlate_brute, x, h)

    def convolve_overlap_save(self, x, h, reverse=0):
        return self._run(self.lib.oracle_convolve_overlap_save, x, h, int(reverse))

    def convolve_fft(self, x, h, reverse=0):
        return self._run(self.lib.oracle_convolve_fft, x, h, int(reverse))

    def convolve(self, x, h):
        return self._run(self.lib.oracle_convolve, x, h)

    def cross_correlate(self, x, h):
        return self._run(self.lib.oracle_cross_correlate, x, h)

    def convolve_f64(self, x, h, reverse=0):
        x, xp = _f32(x)
        h, hp = _f32(h)
        out = np.empty(x.size + h.size - 1, dtype=np.float64)
        self.lib.oracle_convolve_f64(xp, x.size, hp, h.size, int(reverse), out.ctypes.data_as(_dp))
        return out

    def detect_peaks(self, data, type_=3):
        """(positions int32[], values float32[]) of the scalar reference loop."""
        d, dp = _f32(data)
        cap = max(d.size, 1)
        pos = np.empty(cap, dtype=np.int32)
        val = np.empty(cap, dtype=np.float32)
        n = self.lib.oracle_detect_peaks(dp, d.size, int(type_), pos.ctypes.data_as(C.POINTER(C.c_int)),
                                         val.ctypes.data_as(_fp), cap)
        return pos[:n].copy(), val[:n].copy()

    def int16_to_float(self, data):
        d = np.ascontiguousarray(data, dtype=np.int16)
        out = np.empty(d.size, dtype=np.float32)
        self.lib.oracle_int16_to_float(d.ctypes.data, d.size, out.ctypes.data_as(_fp))
        return out

    def float_to_int16(self, data):
        d, dp = _f32(data)
        out = np.empty(d.size, dtype=np.int16)
        self.lib.oracle_float_to_int16(dp, d.size, out.ctypes.data)
        return out

    def minmax1D(self, data):
        d, dp = _f32(data)
        mn, mx = C.c_float(), C.c_float()
        self.lib.oracle_minmax1D(dp, d.size, C.byref(mn), C.byref(mx))
        return mn.value, mx.value

    def overlap_save_length(self, h_length):
        return self.lib.oracle_overlap_save_length(h_length)

    def fft_length(self, x_length, h_length):
        return self.lib.oracle_fft_length(x_length, h_length)

    def select_algorithm(self, x_length, h_length):
        return self.lib.oracle_select_algorithm(x_length, h_length)

    def convolve_batch(self, x, h, reverse=0, threads=0):
        """x: (batch, x_length).  Returns (result (batch, x+h-1), threads_used)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        h, hp = _f32(h)
        b, n = x.shape
        out = np.empty((b, n + h.size - 1), dtype=np.float32)
        used = self.lib.oracle_convolve_batch(x.ctypes.data_as(_fp), n, n, hp, h.size,
                                              out.ctypes.data_as(_fp), out.shape[1], b,
                                              int(reverse), int(threads))
        return out, used


class _OSHandle(C.Structure):
    # /root/reference/inc/simd/convolve_structs.h:39-48
    _fields_ = [("fft_plan", C.c_void_p), ("fft_inverse_plan", C.c_void_p),
                ("fft_boiler_plate", _fp), ("H", _fp), ("x_length", C.c_size_t),
                ("h_length", C.c_size_t), ("L", C.POINTER(C.c_int)), ("reverse", C.c_int)]


class _FFTHandle(C.Structure):
    # /root/reference/inc/simd/convolve_structs.h:50-58
    _fields_ = [("fft_plan", C.c_void_p), ("fft_inverse_plan", C.c_void_p),
                ("M", C.POINTER(C.c_int)), ("x_length", C.c_int), ("h_length", C.c_int),
                ("inputs", C.POINTER(_fp)), ("reverse", C.c_int)]


class _Union(C.Union):
    _fields_ = [("fft", _FFTHandle), ("os", _OSHandle)]


class _Handle(C.Structure):
    # /root/reference/inc/simd/convolve_structs.h:66-74
    _fields_ = [("algorithm", C.c_int), ("x_length", C.c_int), ("h_length", C.c_int),
                ("handle", _Union)]


def bind_simd_api(L):
    """Declare the reference C API (inc/simd/convolve.h, correlate.h) on a ctypes lib.
    Used for oracle/_ref AND for the product libSimd.so -- same 20 symbols."""
    sz = C.c_size_t
    for p in ("convolve", "cross_correlate"):
        getattr(L, p + "_simd").argtypes = [C.c_int, _fp, sz, _fp, sz, _fp]
        getattr(L, p + "_simd").restype = None
        for kind, H in (("_fft", _FFTHandle), ("_overlap_save", _OSHandle), ("", _Handle)):
            ini = getattr(L, p + kind + "_initialize")
            ini.argtypes = [sz, sz]
            ini.restype = H
            run = getattr(L, p + kind)
            run.argtypes = [H, _fp, _fp, _fp]
            run.restype = None
            fin = getattr(L, p + kind + "_finalize")
            fin.argtypes = [H]
            fin.restype = None
    return L


class SimdAPI:
    """Pythonic face of the 20-symbol reference API on any library exporting it."""

    def __init__(self, lib):
        self.lib = bind_simd_api(lib)

    def simd(self, x, h, correlate=False, simd=1):
        x, xp = _f32(x)
        h, hp = _f32(h)
        out = np.empty(x.size + h.size - 1, dtype=np.float32)
        fn = self.lib.cross_correlate_simd if correlate else self.lib.convolve_simd
        fn(int(simd), xp, x.size, hp, h.size, out.ctypes.data_as(_fp))
        return out

    def handle_call(self, kind, x, h, correlate=False, calls=1, return_handle_info=False):
        """kind in {"", "_fft", "_overlap_save"}; runs initialize -> calls x run -> finalize."""
        x, xp = _f32(x)
        h, hp = _f32(h)
        p = "cross_correlate" if correlate else "convolve"
        handle = getattr(self.lib, p + kind + "_initialize")(x.size, h.size)
        outs = []
        for _ in range(calls):
            out = np.full(x.size + h.size - 1, np.nan, dtype=np.float32)
            getattr(self.lib, p + kind)(handle, xp, hp, out.ctypes.data_as(_fp))
            outs.append(out)
        info = None
        if return_handle_info:
            if kind == "":
                info = {"algorithm": handle.algorithm, "x_length": handle.x_length,
                        "h_length": handle.h_length}
                if handle.algorithm == 1:
                    info["M"] = handle.handle.fft.M[0]
                    info["reverse"] = handle.handle.fft.reverse
                elif handle.algorithm == 2:
                    info["L"] = handle.handle.os.L[0]
                    info["reverse"] = handle.handle.os.reverse
            elif kind == "_fft":
                info = {"M": handle.M[0], "reverse": handle.reverse,
                        "x_length": handle.x_length, "h_length": handle.h_length}
            else:
                info = {"L": handle.L[0], "reverse": handle.reverse,
                        "x_length": handle.x_length, "h_length": handle.h_length}
        getattr(self.lib, p + kind + "_finalize")(handle)
        res = outs[0] if calls == 1 else outs
        return (res, info) if return_handle_info else res


class ExtremumPoint(C.Structure):
    # /root/reference/inc/simd/detect_peaks.h:46-49
    _fields_ = [("position", C.c_int), ("value", C.c_float)]


WAVELET_TYPES = {"daubechies": 0, "coiflet": 1, "symlet": 2}   # wavelet_types.h:38-42
EXTENSIONS = {"periodic": 0, "mirror": 1, "constant": 2, "zero": 3}   # wavelet_types.h:44-53


def call_wavelet(lib, type_, order, ext, src, level=None, na=True):
    """wavelet_apply[_na] / stationary_wavelet_apply[_na] through the reference C signatures
    (inc/simd/wavelet.h:81-162) on any library exporting them; host arrays."""
    src, sp = _f32(src)
    n_out = src.size // 2 if level is None else src.size
    hi, lo = np.empty(n_out, np.float32), np.empty(n_out, np.float32)
    name = ("wavelet_apply" if level is None else "stationary_wavelet_apply") + ("_na" if na else "")
    fn = getattr(lib, name)
    fn.restype = None
    if level is None:
        fn.argtypes = [C.c_int, C.c_int, C.c_int, _fp, C.c_size_t, _fp, _fp]
        fn(int(type_), int(order), int(ext), sp, src.size, hi.ctypes.data_as(_fp), lo.ctypes.data_as(_fp))
    else:
        fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, _fp, C.c_size_t, _fp, _fp]
        fn(int(type_), int(order), int(level), int(ext), sp, src.size,
           hi.ctypes.data_as(_fp), lo.ctypes.data_as(_fp))
    return hi, lo


def call_detect_peaks(lib, data, type_=3, simd=0):
    """detect_peaks through the reference C signature on any library exporting it."""
    d, dp = _f32(data)
    lib.detect_peaks.argtypes = [C.c_int, _fp, C.c_size_t, C.c_int,
                                 C.POINTER(C.POINTER(ExtremumPoint)), C.POINTER(C.c_size_t)]
    lib.detect_peaks.restype = None
    res = C.POINTER(ExtremumPoint)()
    n = C.c_size_t(0)
    lib.detect_peaks(int(simd), dp, d.size, int(type_), C.byref(res), C.byref(n))
    if n.value == 0:
        return np.empty(0, np.int32), np.empty(0, np.float32)
    raw = np.ctypeslib.as_array(C.cast(res, C.POINTER(C.c_int32)), shape=(n.value, 2)).copy()
    C.CDLL(None).free(res)
    return raw[:, 0].copy(), raw[:, 1].copy().view(np.float32)


class Reference(SimdAPI):
    """The reference's own convolve.c/correlate.c/memory.c compiled against the FFT shim
    (kind == "reference").  Raises FileNotFoundError when oracle/_ref was not built."""

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO)
        lib = C.CDLL(REF_SO, mode=os.RTLD_LOCAL)
        super().__init__(lib)
        sz = C.c_size_t
        lib.ref_convolve_batch.argtypes = [_fp, sz, sz, _fp, sz, _fp, sz, sz, C.c_int, C.c_int, C.c_int]
        lib.ref_convolve_batch.restype = C.c_int
        lib.ref_max_threads.restype = C.c_int
        lib.ref_struct_layout.argtypes = [C.POINTER(C.c_size_t)]

    def convolve_batch(self, x, h, reverse=0, algorithm=-1, threads=0):
        x = np.ascontiguousarray(x, dtype=np.float32)
        h, hp = _f32(h)
        b, n = x.shape
        out = np.empty((b, n + h.size - 1), dtype=np.float32)
        used = self.lib.ref_convolve_batch(x.ctypes.data_as(_fp), n, n, hp, h.size,
                                           out.ctypes.data_as(_fp), out.shape[1], b,
                                           int(reverse), int(algorithm), int(threads))
        return out, used

    def detect_peaks(self, data, type_=3, simd=0):
        return call_detect_peaks(self.lib, data, type_, simd)

    def int16_to_float(self, data):
        d = np.ascontiguousarray(data, dtype=np.int16)
        out = np.empty(d.size, dtype=np.float32)
        self.lib.ref_int16_to_float_na.argtypes = [C.c_void_p, C.c_size_t, _fp]
        self.lib.ref_int16_to_float_na(d.ctypes.data, d.size, out.ctypes.data_as(_fp))
        return out

    def float_to_int16(self, data):
        d, dp = _f32(data)
        out = np.empty(d.size, dtype=np.int16)
        self.lib.ref_float_to_int16_na.argtypes = [_fp, C.c_size_t, C.c_void_p]
        self.lib.ref_float_to_int16_na(dp, d.size, out.ctypes.data)
        return out

    def minmax1D(self, data, simd=1):
        d, dp = _f32(data)
        self.lib.minmax1D.argtypes = [C.c_int, _fp, C.c_int, _fp, _fp]
        self.lib.minmax1D.restype = None
        mn, mx = C.c_float(), C.c_float()
        self.lib.minmax1D(int(simd), dp, d.size, C.byref(mn), C.byref(mx))
        return mn.value, mx.value

    def wavelet_apply(self, type_, order, ext, src, level=None, na=True):
        return call_wavelet(self.lib, type_, order, ext, src, level, na)

    def wavelet_table(self, type_):
        """The reference's own float table of one family (rows x taps), read from the library."""
        name, rows, width = {0: ("kDaubechiesF", 38, 76), 1: ("kCoifletsF", 5, 30),
                             2: ("kSymletsF", 38, 76)}[int(type_)]
        return np.ctypeslib.as_array((C.c_float * width * rows).in_dll(self.lib, name)).copy()

    def struct_layout(self):
        arr = (C.c_size_t * 16)()
        self.lib.ref_struct_layout(arr)
        return list(arr)


def have_reference():
    return os.path.exists(REF_SO)


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / (den if den > 0 else 1.0))


def parity_ok(got, want, x, h, rel=1e-5, abs_k=1e-4):
    """BASELINE.json north_star tolerance: rel-L2 <= 1e-5 and
    max|got-want| <= 1e-4 * ||h||_1 * max|x|."""
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    if got.shape != want.shape or not np.all(np.isfinite(got)):
        return False, float("inf"), float("inf")
    r = rel_l2(got, want)
    m = float(np.max(np.abs(got - want))) if got.size else 0.0
    bound = abs_k * float(np.sum(np.abs(np.asarray(h, dtype=np.float64)))) * float(np.max(np.abs(x)))
    return (r <= rel and m <= bound), r, m
