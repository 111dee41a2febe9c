"""veles.simd_b200 -- Python face of libSimd.so (B200 build of the veles.simd convolve /
cross-correlate hot path).

This module is a thin ctypes binding over the C-ABI library built from src/*.c and
csrc/*.cu; it exists for tests, the benchmark and Python callers.  The product is the
shared library: C callers link it directly through include/simd/*.h.

There is NO CPU fallback.  Importing works anywhere (so that host-side logic such as the
selector can be exercised without a GPU), but every compute call goes to the CUDA kernels
and aborts the process with a message if no CUDA device is usable; a missing libSimd.so
raises ImportError at import time.

Names mirror the reference API (/root/reference/inc/simd/convolve.h, correlate.h):
    lib.convolve_simd, lib.convolve_initialize, ...   raw ctypes functions
    convolve(x, h), cross_correlate(x, h)             numpy / torch convenience wrappers
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# VELES_B200_LIB: another build of the same library (A/B timing of two builds on one GPU box)
LIB_PATH = os.environ.get("VELES_B200_LIB") or os.path.join(_HERE, "libSimd.so")

_fp = C.POINTER(C.c_float)

ALG_BRUTE_FORCE, ALG_FFT, ALG_OVERLAP_SAVE = 0, 1, 2


class ConvolutionOverlapSaveHandle(C.Structure):
    """include/simd/convolve_structs.h (reference: convolve_structs.h:39-48)."""
    _fields_ = [("fft_plan", C.c_void_p), ("fft_inverse_plan", C.c_void_p),
                ("fft_boiler_plate", C.c_void_p), ("H", C.c_void_p),
                ("x_length", C.c_size_t), ("h_length", C.c_size_t),
                ("L", C.POINTER(C.c_int)), ("reverse", C.c_int)]


class ConvolutionFFTHandle(C.Structure):
    """include/simd/convolve_structs.h (reference: convolve_structs.h:50-58)."""
    _fields_ = [("fft_plan", C.c_void_p), ("fft_inverse_plan", C.c_void_p),
                ("M", C.POINTER(C.c_int)), ("x_length", C.c_int), ("h_length", C.c_int),
                ("inputs", C.c_void_p), ("reverse", C.c_int)]


class _HandleUnion(C.Union):
    _fields_ = [("fft", ConvolutionFFTHandle), ("os", ConvolutionOverlapSaveHandle)]


class ConvolutionHandle(C.Structure):
    """include/simd/convolve_structs.h (reference: convolve_structs.h:66-74)."""
    _fields_ = [("algorithm", C.c_int), ("x_length", C.c_int), ("h_length", C.c_int),
                ("handle", _HandleUnion)]


class ExtremumPoint(C.Structure):
    """include/simd/detect_peaks.h (reference: detect_peaks.h:46-49)."""
    _fields_ = [("position", C.c_int), ("value", C.c_float)]


EXTREMUM_MAXIMUM, EXTREMUM_MINIMUM, EXTREMUM_BOTH = 1, 2, 3


def build(verbose=False):
    """Compile libSimd.so in-tree (nvcc -gencode arch=compute_100a,code=sm_100a)."""
    subprocess.run(["make", "-C", _HERE, "--no-print-directory"] + ([] if verbose else ["-s"]),
                   check=True)


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C {_HERE}` "
            "(or __graft_entry__.build()); veles.simd_b200 has no CPU fallback")
    L = C.CDLL(LIB_PATH)
    sz, vp = C.c_size_t, C.c_void_p
    for p in ("convolve", "cross_correlate"):
        f = getattr(L, p + "_simd")
        f.argtypes = [C.c_int, vp, sz, vp, sz, vp]
        f.restype = None
        for kind, H in (("_fft", ConvolutionFFTHandle),
                        ("_overlap_save", ConvolutionOverlapSaveHandle),
                        ("", ConvolutionHandle)):
            ini = getattr(L, p + kind + "_initialize")
            ini.argtypes = [sz, sz]
            ini.restype = H
            run = getattr(L, p + kind)
            run.argtypes = [H, vp, vp, vp]
            run.restype = None
            fin = getattr(L, p + kind + "_finalize")
            fin.argtypes = [H]
            fin.restype = None
        b = getattr(L, p + "_batch")
        b.argtypes = [ConvolutionHandle, vp, sz, vp, vp, sz, sz]
        b.restype = None
    for p in ("convolve", "cross_correlate"):
        b = getattr(L, p + "_batch_int16")
        b.argtypes = [ConvolutionHandle, vp, sz, vp, vp, sz, sz]
        b.restype = None
    L.simd_b200_cross_correlate_peaks.argtypes = [ConvolutionHandle, vp, vp, C.c_int,
                                                  C.POINTER(C.POINTER(ExtremumPoint)), C.POINTER(sz)]
    L.simd_b200_cross_correlate_peaks.restype = None
    L.simd_b200_filter_shard.argtypes = [vp, sz, vp, sz, C.c_int, sz, vp, sz, C.c_int]
    L.simd_b200_filter_shard.restype = None
    L.simd_b200_mallocf_pinned.argtypes = [sz]
    L.simd_b200_mallocf_pinned.restype = vp
    L.simd_b200_free_pinned.argtypes = [vp]
    L.simd_b200_free_pinned.restype = None
    L.simd_b200_use_devices.argtypes = [C.c_int]
    L.simd_b200_use_devices.restype = None
    L.simd_b200_devices_in_use.restype = C.c_int
    L.simd_b200_filter_long.argtypes = [vp, sz, vp, sz, C.c_int, vp]
    L.simd_b200_filter_long.restype = None
    L.simd_b200_filter_long_sharded.argtypes = [C.c_int, C.POINTER(C.c_int), C.POINTER(vp), C.POINTER(sz),
                                                C.POINTER(vp), sz, C.c_int, C.POINTER(vp)]
    L.simd_b200_filter_long_sharded.restype = None
    L.simd_b200_select_algorithm.argtypes = [sz, sz]
    L.simd_b200_select_algorithm.restype = C.c_int
    L.simd_b200_select_algorithm_batch.argtypes = [sz, sz, sz]
    L.simd_b200_select_algorithm_batch.restype = C.c_int
    L.simd_b200_transform_length.argtypes = [sz, sz]
    L.simd_b200_transform_length.restype = C.c_int
    L.simd_b200_set_stream.argtypes = [vp]
    L.simd_b200_set_stream.restype = None
    L.simd_b200_get_stream.restype = vp
    L.simd_b200_synchronize.restype = None
    L.simd_b200_device_count.restype = C.c_int
    L.simd_b200_kernel_launches.restype = C.c_longlong
    L.simd_b200_profile.argtypes = [C.c_int]
    L.simd_b200_profile.restype = None
    L.simd_b200_profile_collect.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
    L.simd_b200_profile_collect.restype = None
    L.detect_peaks.argtypes = [C.c_int, vp, sz, C.c_int, C.POINTER(C.POINTER(ExtremumPoint)),
                               C.POINTER(sz)]
    L.detect_peaks.restype = None
    L.simd_b200_detect_peaks_device.argtypes = [vp, sz, C.c_int, vp, sz]
    L.simd_b200_detect_peaks_device.restype = sz
    L.minmax1D.argtypes = [C.c_int, vp, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.minmax1D.restype = None
    L.simd_b200_minmax1D_device.argtypes = [vp, sz, vp]
    L.simd_b200_minmax1D_device.restype = None
    L.int16_to_float.argtypes = [vp, sz, vp]
    L.int16_to_float.restype = None
    L.float_to_int16.argtypes = [vp, sz, vp]
    L.float_to_int16.restype = None
    L.simd_b200_filter_noalloc.argtypes = [vp, sz, vp, sz, C.c_int, vp, vp]
    L.simd_b200_filter_noalloc.restype = C.c_int
    L.wavelet_validate_order.argtypes = [C.c_int, C.c_int]
    L.wavelet_validate_order.restype = C.c_int
    L.wavelet_prepare_array.argtypes = [C.c_int, vp, sz]
    L.wavelet_prepare_array.restype = vp
    L.wavelet_allocate_destination.argtypes = [C.c_int, sz]
    L.wavelet_allocate_destination.restype = vp
    L.wavelet_recycle_source.argtypes = [C.c_int, vp, sz] + [C.POINTER(vp)] * 4
    L.wavelet_recycle_source.restype = None
    for name in ("wavelet_apply", "wavelet_apply_na"):
        getattr(L, name).argtypes = [C.c_int, C.c_int, C.c_int, vp, sz, vp, vp]
        getattr(L, name).restype = None
    for name in ("stationary_wavelet_apply", "stationary_wavelet_apply_na"):
        getattr(L, name).argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, vp, sz, vp, vp]
        getattr(L, name).restype = None
    L.simd_b200_wavelet_filters.argtypes = [C.c_int, C.c_int, vp, vp]
    L.simd_b200_wavelet_filters.restype = C.c_int
    L.mallocf.argtypes = [sz]
    L.mallocf.restype = vp
    L.malloc_aligned.argtypes = [sz]
    L.malloc_aligned.restype = vp
    L.memsetf.argtypes = [vp, C.c_float, sz]
    L.memsetf.restype = None
    L.rmemcpyf.argtypes = [vp, vp, sz]
    L.rmemcpyf.restype = vp
    return L


lib = _load()

#: every symbol include/simd/*.h declares (checked by tests/test_abi.py)
EXPORTED_SYMBOLS = (
    "convolve_simd convolve_overlap_save_initialize convolve_overlap_save "
    "convolve_overlap_save_finalize convolve_fft_initialize convolve_fft convolve_fft_finalize "
    "convolve_initialize convolve convolve_finalize cross_correlate_simd "
    "cross_correlate_fft_initialize cross_correlate_fft cross_correlate_fft_finalize "
    "cross_correlate_overlap_save_initialize cross_correlate_overlap_save "
    "cross_correlate_overlap_save_finalize cross_correlate_initialize cross_correlate "
    "cross_correlate_finalize mallocf malloc_aligned memsetf rmemcpyf convolve_batch "
    "cross_correlate_batch simd_b200_filter_shard simd_b200_select_algorithm "
    "simd_b200_select_algorithm_batch simd_b200_mallocf_pinned simd_b200_free_pinned simd_b200_use_devices simd_b200_devices_in_use "
    "simd_b200_filter_long simd_b200_filter_long_sharded convolve_batch_int16 "
    "cross_correlate_batch_int16 simd_b200_cross_correlate_peaks "
    "simd_b200_transform_length simd_b200_set_stream simd_b200_get_stream "
    "simd_b200_synchronize simd_b200_device_count simd_b200_kernel_launches "
    "simd_b200_profile simd_b200_profile_collect simd_b200_filter_noalloc "
    "detect_peaks simd_b200_detect_peaks_device minmax1D simd_b200_minmax1D_device "
    "int16_to_float float_to_int16 wavelet_validate_order wavelet_prepare_array "
    "wavelet_allocate_destination wavelet_recycle_source wavelet_apply wavelet_apply_na "
    "stationary_wavelet_apply stationary_wavelet_apply_na simd_b200_wavelet_filters").split()


def _is_torch(a):
    return type(a).__module__.startswith("torch")


_installed_stream = {}     # device index -> stream handle last given to the library


def _follow_torch_stream(device_index):
    """The library enqueues on its own non-blocking stream unless told otherwise.  A CUDA
    tensor's producer (a torch kernel, an NCCL receive that req.wait() ordered) lives on
    torch's current stream, so work on CUDA tensors must go to that same stream or it can
    read data that has not landed yet: install it before every such call."""
    import torch
    if torch.cuda.current_device() != device_index:
        raise ValueError(f"tensor on cuda:{device_index} but cuda:{torch.cuda.current_device()} is current "
                         "(the library works on the calling thread's current device)")
    handle = torch.cuda.current_stream(device_index).cuda_stream
    if _installed_stream.get(device_index) != handle:
        lib.simd_b200_set_stream(C.c_void_p(handle))
        _installed_stream[device_index] = handle


def _ptr(a):
    """Raw address of a numpy array (host) or torch tensor (host or device)."""
    if _is_torch(a):
        # 1-D contiguous, or 2-D with contiguous rows (row stride is passed separately)
        assert a.dtype.is_floating_point and a.element_size() == 4 and a.stride(-1) == 1
        if a.is_cuda:
            _follow_torch_stream(a.device.index)
        return a.data_ptr()
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


def _empty_like(a, shape):
    if _is_torch(a):
        import torch
        return torch.empty(shape, dtype=torch.float32, device=a.device)
    return np.empty(shape, dtype=np.float32)


def use_torch_stream():
    """Make the library enqueue on torch's current CUDA stream (so torch.cuda.Event timing
    and torch allocator stream-ordering see the kernels).  The wrappers in this module do
    this on their own whenever they are handed a CUDA tensor; raw lib.* calls with
    data_ptr() arguments need it once per stream change."""
    import torch
    _follow_torch_stream(torch.cuda.current_device())


def profile(enable):
    lib.simd_b200_profile(int(enable))


def profile_collect():
    """{kind: (total_ms, launches)} for kinds direct / fused / spectrum since the last call."""
    ms = (C.c_double * 3)()
    cnt = (C.c_longlong * 3)()
    lib.simd_b200_profile_collect(ms, cnt)
    return {k: (ms[i], cnt[i]) for i, k in enumerate(("direct", "fused", "spectrum"))}


def select_algorithm(x_length, h_length, batch=1):
    return lib.simd_b200_select_algorithm_batch(x_length, h_length, batch)


def transform_length(x_length, h_length):
    return lib.simd_b200_transform_length(x_length, h_length)


def _prep(x, h):
    if not _is_torch(x):
        x = np.ascontiguousarray(x, dtype=np.float32)
    if not _is_torch(h):
        h = np.ascontiguousarray(h, dtype=np.float32)
    return x, h


def convolve_simd(x, h, correlate=False, simd=1, out=None):
    """convolve_simd / cross_correlate_simd (direct kernel).  1-D x and h."""
    x, h = _prep(x, h)
    n, m = x.shape[-1], h.shape[-1]
    if out is None:
        out = _empty_like(x, (n + m - 1,))
    fn = lib.cross_correlate_simd if correlate else lib.convolve_simd
    fn(int(simd), _ptr(x), n, _ptr(h), m, _ptr(out))
    return out


class Handle:
    """RAII wrapper over the three handle families.
    kind: "" (best method), "_fft", "_overlap_save"; correlate selects cross_correlate_*."""

    def __init__(self, x_length, h_length, kind="", correlate=False):
        self.kind, self.correlate = kind, correlate
        self.prefix = "cross_correlate" if correlate else "convolve"
        self.x_length, self.h_length = int(x_length), int(h_length)
        self.handle = getattr(lib, self.prefix + kind + "_initialize")(self.x_length, self.h_length)
        self._open = True

    @property
    def algorithm(self):
        if self.kind == "":
            return self.handle.algorithm
        return ALG_FFT if self.kind == "_fft" else ALG_OVERLAP_SAVE

    @property
    def transform_length(self):
        """L (overlap-save) or M (FFT) as the handle reports it; None for brute force."""
        if self.kind == "_fft":
            return self.handle.M[0]
        if self.kind == "_overlap_save":
            return self.handle.L[0]
        if self.handle.algorithm == ALG_FFT:
            return self.handle.handle.fft.M[0]
        if self.handle.algorithm == ALG_OVERLAP_SAVE:
            return self.handle.handle.os.L[0]
        return None

    @property
    def reverse(self):
        if self.kind != "":
            return self.handle.reverse
        if self.handle.algorithm == ALG_FFT:
            return self.handle.handle.fft.reverse
        if self.handle.algorithm == ALG_OVERLAP_SAVE:
            return self.handle.handle.os.reverse
        return None

    def __call__(self, x, h, out=None):
        """One signal (1-D x) through <prefix><kind>(), or a batch (2-D x, best-method
        handles only) through <prefix>_batch()."""
        x, h = _prep(x, h)
        assert x.shape[-1] == self.x_length and h.shape[-1] == self.h_length
        olen = self.x_length + self.h_length - 1
        if x.ndim == 1:
            if out is None:
                out = _empty_like(x, (olen,))
            getattr(lib, self.prefix + self.kind)(self.handle, _ptr(x), _ptr(h), _ptr(out))
            return out
        assert x.ndim == 2 and self.kind == "", "batches go through the best-method handle"
        if out is None:
            out = _empty_like(x, (x.shape[0], olen))
        x_stride = x.stride(0) if _is_torch(x) else x.strides[0] // 4
        o_stride = out.stride(0) if _is_torch(out) else out.strides[0] // 4
        getattr(lib, self.prefix + "_batch")(self.handle, _ptr(x), x_stride, _ptr(h), _ptr(out),
                                             o_stride, x.shape[0])
        return out

    def close(self):
        if self._open:
            getattr(lib, self.prefix + self.kind + "_finalize")(self.handle)
            self._open = False

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def convolve(x, h, correlate=False):
    """initialize -> run -> finalize through the best-method handle.  x may be 1-D (one
    signal) or 2-D (a batch of independent signals sharing h)."""
    with Handle(x.shape[-1], h.shape[-1], "", correlate) as hd:
        out = hd(x, h)
        if _is_torch(out) and out.is_cuda:
            lib.simd_b200_synchronize()
    return out


def cross_correlate(x, h):
    return convolve(x, h, correlate=True)


def minmax1D(data, simd=1):
    """minmax1D(): (min, max) of a host or device series."""
    if not _is_torch(data):
        data = np.ascontiguousarray(data, dtype=np.float32)
    mn, mx = C.c_float(), C.c_float()
    lib.minmax1D(int(simd), _ptr(data), data.shape[-1], C.byref(mn), C.byref(mx))
    return mn.value, mx.value


#: WaveletType / ExtensionType (include/simd/wavelet_types.h)
WAVELET_DAUBECHIES, WAVELET_COIFLET, WAVELET_SYMLET = 0, 1, 2
EXTENSION_PERIODIC, EXTENSION_MIRROR, EXTENSION_CONSTANT, EXTENSION_ZERO = 0, 1, 2, 3


def wavelet_filters(type_, order):
    """simd_b200_wavelet_filters: (highpass, lowpass) float32 arrays, or None when (type,
    order) has no table row."""
    if order < 1:
        return None
    hp, lp = np.empty(order, np.float32), np.empty(order, np.float32)
    if not lib.simd_b200_wavelet_filters(int(type_), int(order), hp.ctypes.data, lp.ctypes.data):
        return None
    return hp, lp


def wavelet_apply(type_, order, ext, src, level=None, na=False):
    """wavelet_apply[_na] (level None) or stationary_wavelet_apply[_na]: (desthi, destlo) of a
    host (numpy) or device (torch CUDA) series; results live where the input lives."""
    if not _is_torch(src):
        src = np.ascontiguousarray(src, dtype=np.float32)
    n = src.shape[-1]
    n_out = n // 2 if level is None else n
    hi, lo = _empty_like(src, (n_out,)), _empty_like(src, (n_out,))
    if level is None:
        fn = lib.wavelet_apply_na if na else lib.wavelet_apply
        fn(int(type_), int(order), int(ext), _ptr(src), n, _ptr(hi), _ptr(lo))
    else:
        fn = lib.stationary_wavelet_apply_na if na else lib.stationary_wavelet_apply
        fn(int(type_), int(order), int(level), int(ext), _ptr(src), n, _ptr(hi), _ptr(lo))
    return hi, lo


def detect_peaks(data, type_=EXTREMUM_BOTH, simd=1):
    """detect_peaks(): (positions int32[], values float32[]) of a host or device series."""
    if not _is_torch(data):
        data = np.ascontiguousarray(data, dtype=np.float32)
    res = C.POINTER(ExtremumPoint)()
    n = C.c_size_t(0)
    lib.detect_peaks(int(simd), _ptr(data), data.shape[-1], int(type_), C.byref(res), C.byref(n))
    if n.value == 0:
        return np.empty(0, np.int32), np.empty(0, np.float32)
    raw = np.ctypeslib.as_array(C.cast(res, C.POINTER(C.c_int32)), shape=(n.value, 2)).copy()
    C.CDLL(None).free(res)
    return raw[:, 0].copy(), raw[:, 1].copy().view(np.float32)


def detect_peaks_device(data, out, type_=EXTREMUM_BOTH):
    """simd_b200_detect_peaks_device: data, out are CUDA tensors (out: int32 (capacity, 2));
    returns the number of extrema found (may exceed the capacity)."""
    return lib.simd_b200_detect_peaks_device(data.data_ptr(), data.numel(), int(type_), out.data_ptr(),
                                             out.shape[0])


def filter_shard(x, h, reverse, left_pad, result_length, algorithm=-1, out=None):
    """simd_b200_filter_shard: result[n] = sum_j g[j] X(n + j - left_pad)."""
    x, h = _prep(x, h)
    if out is None:
        out = _empty_like(x, (result_length,))
    lib.simd_b200_filter_shard(_ptr(x), x.shape[-1], _ptr(h), h.shape[-1], int(reverse),
                               int(left_pad), _ptr(out), int(result_length), int(algorithm))
    return out


def batch_int16(handle, x16, h, out=None):
    """convolve_batch_int16 / cross_correlate_batch_int16: x16 is a host int16 array (batch, x_len)
    (numpy, or a pinned torch tensor); h and out are float32."""
    assert handle.kind == ""
    if _is_torch(x16):
        assert x16.dtype.itemsize == 2 and not x16.is_cuda and x16.stride(-1) == 1
        xp, xs, nb = x16.data_ptr(), x16.stride(0), x16.shape[0]
    else:
        x16 = np.ascontiguousarray(x16, dtype=np.int16)
        xp, xs, nb = x16.ctypes.data, x16.strides[0] // 2, x16.shape[0]
    olen = handle.x_length + handle.h_length - 1
    if out is None:
        out = np.empty((nb, olen), dtype=np.float32)
    os_ = out.stride(0) if _is_torch(out) else out.strides[0] // 4
    if not _is_torch(h):
        h = np.ascontiguousarray(h, dtype=np.float32)
    getattr(lib, handle.prefix + "_batch_int16")(handle.handle, xp, xs, _ptr(h), _ptr(out), os_, nb)
    return out


def cross_correlate_peaks(x, h, type_=EXTREMUM_BOTH):
    """simd_b200_cross_correlate_peaks: (positions, values) of the extrema of x (x) h; the
    correlation itself stays on the GPU."""
    x, h = _prep(x, h)
    res = C.POINTER(ExtremumPoint)()
    n = C.c_size_t(0)
    with Handle(x.shape[-1], h.shape[-1], "", True) as hd:
        lib.simd_b200_cross_correlate_peaks(hd.handle, _ptr(x), _ptr(h), int(type_), C.byref(res), C.byref(n))
    if n.value == 0:
        return np.empty(0, np.int32), np.empty(0, np.float32)
    raw = np.ctypeslib.as_array(C.cast(res, C.POINTER(C.c_int32)), shape=(n.value, 2)).copy()
    C.CDLL(None).free(res)
    return raw[:, 0].copy(), raw[:, 1].copy().view(np.float32)


class PinnedArray:
    """numpy view of simd_b200_mallocf_pinned memory (freed with the object)."""

    def __init__(self, shape):
        n = int(np.prod(shape))
        self.ptr = lib.simd_b200_mallocf_pinned(n)
        self.array = np.ctypeslib.as_array(C.cast(self.ptr, C.POINTER(C.c_float)), shape=(n,)).reshape(shape)

    def close(self):
        if self.ptr:
            self.array = None
            lib.simd_b200_free_pinned(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def use_devices(n):
    """simd_b200_use_devices: host-pointer batches / long signals spread over n GPUs (0 = all)."""
    lib.simd_b200_use_devices(int(n))
    return lib.simd_b200_devices_in_use()


def filter_long(x, h, reverse=False, out=None):
    """simd_b200_filter_long: full convolution / correlation of one long signal (host arrays
    are spread over the GPUs chosen with use_devices)."""
    x, h = _prep(x, h)
    if out is None:
        out = _empty_like(x, (x.shape[-1] + h.shape[-1] - 1,))
    lib.simd_b200_filter_long(_ptr(x), x.shape[-1], _ptr(h), h.shape[-1], int(bool(reverse)), _ptr(out))
    return out


def filter_long_sharded(x_shards, h_copies, reverse=False, results=None):
    """simd_b200_filter_long_sharded over torch CUDA tensors, one shard per device.  Every
    x_shards[i] must be a view that has h-1 elements of storage in front of it (see
    sharded.alloc_shard); h_copies[i] is h on x_shards[i]'s device."""
    import torch
    n = len(x_shards)
    hl = h_copies[0].numel()
    if results is None:
        results = [torch.empty(x.numel() + (hl - 1 if i == n - 1 else 0), device=x.device)
                   for i, x in enumerate(x_shards)]
    for x in x_shards:                     # each device's library stream follows torch's
        with torch.cuda.device(x.device):
            _follow_torch_stream(x.device.index)
    devs = (C.c_int * n)(*[x.device.index for x in x_shards])
    xs = (C.c_void_p * n)(*[x.data_ptr() for x in x_shards])
    lens = (C.c_size_t * n)(*[x.numel() for x in x_shards])
    hs = (C.c_void_p * n)(*[h.data_ptr() for h in h_copies])
    rs = (C.c_void_p * n)(*[r.data_ptr() for r in results])
    lib.simd_b200_filter_long_sharded(n, devs, xs, lens, hs, hl, int(bool(reverse)), rs)
    return results
