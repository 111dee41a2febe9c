"""Error behaviour of the drop-in library, checked in child processes because the reference
convention -- kept here -- is to abort the process on a violated precondition
(/root/reference/src/convolve.c:44-48,105-107: assert(), live in release builds)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_child(code):
    from conftest import _ensure_built
    _ensure_built()
    env = dict(os.environ, PYTHONPATH=ROOT)
    return subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env,
                          timeout=300)


def test_overlap_save_initialize_rejects_long_filters():
    # reference: assert(hLength < xLength / 2), src/convolve.c:105 -- host-side check, no GPU needed
    r = run_child("import veles.simd_b200 as v; v.lib.convolve_overlap_save_initialize(10, 8)")
    assert r.returncode != 0
    assert "precondition failed" in r.stderr and "hLength < xLength / 2" in r.stderr


def test_zero_lengths_abort():
    r = run_child("import veles.simd_b200 as v; v.lib.convolve_initialize(0, 4)")
    assert r.returncode != 0 and "precondition failed" in r.stderr


def test_no_cpu_fallback_without_a_gpu():
    """On a machine without a CUDA device every compute entry point must fail loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for machines without a GPU")
    r = run_child("import numpy as np, veles.simd_b200 as v\n"
                  "x = np.ones(100, np.float32); h = np.ones(5, np.float32)\n"
                  "print(v.convolve_simd(x, h)[:3])")
    assert r.returncode != 0
    assert "no CPU fallback" in r.stderr or "CUDA" in r.stderr
    assert "[" not in r.stdout            # nothing was computed


@pytest.mark.gpu
def test_null_pointer_aborts_on_gpu_box():
    r = run_child("import veles.simd_b200 as v; v.lib.convolve_simd(1, None, 8, None, 4, None)")
    assert r.returncode != 0 and "precondition failed" in r.stderr


@pytest.mark.gpu
def test_handles_and_handleless_calls_from_several_threads():
    """Threading contract: distinct handles may be used concurrently, the
    handle-less *_simd calls are reentrant."""
    code = r'''
import threading, numpy as np, sys
sys.path.insert(0, "tests")
import veles.simd_b200 as v
from _oracle import Oracle, parity_ok
o = Oracle()
errors = []
def work(seed):
    rng = np.random.default_rng(seed)
    xl, hl = 30000 + 1000 * seed, 200 + 50 * seed
    x = rng.uniform(-1, 1, xl).astype(np.float32); h = rng.uniform(-1, 1, hl).astype(np.float32)
    want = o.convolve_f64(x, h)
    for kind in ("_overlap_save", "_fft", ""):
        with v.Handle(xl, hl, kind) as hd:
            for _ in range(3):
                ok, r, m = parity_ok(hd(x, h), want, x, h)
                if not ok: errors.append((seed, kind, r, m))
    for _ in range(3):
        ok, r, m = parity_ok(v.convolve_simd(x[:5000], h[:64]), o.convolve_f64(x[:5000], h[:64]), x, h)
        if not ok: errors.append((seed, "simd", r, m))
ts = [threading.Thread(target=work, args=(s,)) for s in range(6)]
[t.start() for t in ts]; [t.join() for t in ts]
print("ERRORS", errors)
assert not errors
'''
    r = run_child(code)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "ERRORS []" in r.stdout
