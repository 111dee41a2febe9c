"""CPU tests of the drop-in boundary: the C-ABI library loads without a GPU, exports every
symbol include/simd/*.h declares, keeps the reference's struct layout, and its host-side
logic (selector, transform-length policy, memory helpers) behaves."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")


def _declared_functions():
    names = set()
    for fn in sorted(os.listdir(os.path.join(INC, "simd"))):
        src = open(os.path.join(INC, "simd", fn)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        src = re.sub(r"#.*", "", src)
        for m in re.finditer(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\(", src):
            n = m.group(1)
            if n.startswith(("convolve", "cross_correlate", "simd_b200", "malloc", "memsetf",
                             "rmemcpyf", "detect_peaks", "minmax1D", "int16_to_float", "float_to_int16",
                             "wavelet_", "stationary_wavelet_")):
                names.add(n)
    return names


def test_library_exports_every_declared_symbol(simd):
    declared = _declared_functions()
    assert len(declared) >= 51
    out = subprocess.run(["nm", "-D", "--defined-only", simd.LIB_PATH], capture_output=True,
                         text=True, check=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    missing = declared - exported
    assert not missing, missing
    assert declared == set(simd.EXPORTED_SYMBOLS)
    for n in declared:
        assert getattr(simd.lib, n) is not None


def test_no_oracle_or_cpu_fallback_linked(simd):
    out = subprocess.run(["nm", "-D", simd.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle_" not in out and "fftf_" not in out and "ref_convolve" not in out
    deps = subprocess.run(["ldd", simd.LIB_PATH], capture_output=True, text=True).stdout
    assert "liboracle" not in deps and "libSimd_ref" not in deps and "cufft" not in deps.lower()


def test_struct_layout_matches_reference_abi(simd):
    # SURVEY 8(b) [probed on x86-64 gcc 13]: 64 / 48 / 80 bytes and these offsets.
    O, F, H = simd.ConvolutionOverlapSaveHandle, simd.ConvolutionFFTHandle, simd.ConvolutionHandle
    assert (C.sizeof(O), C.sizeof(F), C.sizeof(H)) == (64, 48, 80)
    assert [getattr(O, f).offset for f in ("fft_plan", "fft_inverse_plan", "fft_boiler_plate", "H",
                                           "x_length", "h_length", "L", "reverse")] == \
        [0, 8, 16, 24, 32, 40, 48, 56]
    assert [getattr(F, f).offset for f in ("fft_plan", "fft_inverse_plan", "M", "x_length",
                                           "h_length", "inputs", "reverse")] == [0, 8, 16, 24, 28, 32, 40]
    assert [getattr(H, f).offset for f in ("algorithm", "x_length", "h_length", "handle")] == [0, 4, 8, 16]


def test_struct_layout_compiled_from_headers(tmp_path, reference):
    """The C compiler's view of include/simd/*.h equals the reference headers' layout."""
    src = tmp_path / "layout.c"
    src.write_text(r'''
#include <stdio.h>
#include <stddef.h>
#include <simd/convolve.h>
#include <simd/correlate.h>
#include <simd/b200.h>
int main(void) {
  printf("%zu %zu %zu ", sizeof(ConvolutionOverlapSaveHandle), sizeof(ConvolutionFFTHandle), sizeof(ConvolutionHandle));
  printf("%zu %zu %zu %zu %zu %zu ", offsetof(ConvolutionOverlapSaveHandle, fft_boiler_plate), offsetof(ConvolutionOverlapSaveHandle, H),
         offsetof(ConvolutionOverlapSaveHandle, x_length), offsetof(ConvolutionOverlapSaveHandle, h_length),
         offsetof(ConvolutionOverlapSaveHandle, L), offsetof(ConvolutionOverlapSaveHandle, reverse));
  printf("%zu %zu %zu %zu %zu ", offsetof(ConvolutionFFTHandle, M), offsetof(ConvolutionFFTHandle, x_length),
         offsetof(ConvolutionFFTHandle, h_length), offsetof(ConvolutionFFTHandle, inputs), offsetof(ConvolutionFFTHandle, reverse));
  printf("%zu %zu\n", offsetof(ConvolutionHandle, x_length), offsetof(ConvolutionHandle, handle));
  return kConvolutionAlgorithmBruteForce != 0 || kConvolutionAlgorithmFFT != 1 || kConvolutionAlgorithmOverlapSave != 2;
}''')
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c99", "-I", INC, str(src), "-o", str(exe)], check=True)
    mine = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True,
                                            check=True).stdout.split()]
    assert mine == reference.struct_layout()


def test_headers_compile_as_cxx(tmp_path):
    src = tmp_path / "t.cc"
    src.write_text('#include <simd/convolve.h>\n#include <simd/correlate.h>\n#include <simd/memory.h>\n'
                   '#include <simd/b200.h>\nint main() { return 0; }\n')
    subprocess.run(["g++", "-std=c++11", "-fsyntax-only", "-I", INC, str(src)], check=True)


def test_selector_is_pure_host_logic(simd):
    # BASELINE configs on the measured B200 cost model (profiles/selector_sweep_r01.md):
    # single small signals are launch-bound -> one direct launch; C4/C5 -> overlap-save;
    # the C3 *batch* (4096 signals, h=64) crosses over to the transform path
    assert simd.select_algorithm(1024, 50) == simd.ALG_BRUTE_FORCE
    assert simd.select_algorithm(16384, 64) == simd.ALG_BRUTE_FORCE
    assert simd.select_algorithm(16384, 64, batch=4096) == simd.ALG_OVERLAP_SAVE
    assert simd.select_algorithm(16384, 8, batch=4096) == simd.ALG_BRUTE_FORCE
    assert simd.select_algorithm(65536, 512) == simd.ALG_BRUTE_FORCE
    assert simd.select_algorithm(1 << 20, 512) == simd.ALG_OVERLAP_SAVE
    assert simd.select_algorithm(262144, 4096) == simd.ALG_OVERLAP_SAVE
    assert simd.select_algorithm(2 ** 30, 8192) == simd.ALG_OVERLAP_SAVE
    assert simd.select_algorithm(100000, 90000) == simd.ALG_FFT        # x <= 2h, large
    for xl, hl in [(65536, 512), (262144, 4096), (2 ** 30, 8192), (100000, 90000), (5000, 3000)]:
        n = simd.transform_length(xl, hl)
        assert 256 <= n <= 16384 and n & (n - 1) == 0
    assert simd.transform_length(5000, 3000) == 8192      # whole signal: next pow2 >= x+h-1
    assert simd.transform_length(262144, 4096) == 16384
    assert simd.transform_length(1 << 20, 64) == 256       # two-pass plan, 8 transforms per CTA
    assert simd.transform_length(1 << 20, 300) == 2048


def test_memory_helpers(simd):
    p = simd.lib.mallocf(100)
    assert p and p % 64 == 0
    a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(100,))
    simd.lib.memsetf(p, 2.5, 100)
    assert np.all(a == 2.5)
    src = np.arange(100, dtype=np.float32)
    simd.lib.rmemcpyf(p, src.ctypes.data, 100)
    assert np.array_equal(a, src[::-1])
    C.CDLL(None).free(C.c_void_p(p))


def test_emulated_kernel_logic(tmp_path):
    """The fused overlap-save kernel's index / twiddle logic (fft_core.cuh), executed on the
    CPU thread by thread for every plan, against a float64 direct convolution."""
    exe = tmp_path / "emul_os"
    subprocess.run(["g++", "-std=c++17", "-O1", "-o", str(exe),
                    os.path.join(ROOT, "tests", "emul", "emul_os.cpp")], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0 and "EMUL OK" in r.stdout, r.stdout
