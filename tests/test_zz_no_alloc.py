"""GPU test of simd_b200_filter_noalloc, the entry point for callers that cannot allocate
device memory while they run (fixed-budget arenas, plug-in hosts that forbid cudaMalloc
after start-up, real-time loops).  It must give the same numbers as the ordinary handle
path, which allocates its scratch in *_initialize, and both must agree with the CPU oracle.

Kept in a file that pytest collects last, so that a failure here cannot hide the result of
any other test in a run that stops at the first failure.  Called through the C-ABI (ctypes)
with device pointers, twice in a row with different data, because every call reuses the
same built-in scratch array and must not depend on what the previous call left there.
"""
import ctypes as C

import numpy as np
import pytest

from _oracle import parity_ok, rel_l2

pytestmark = pytest.mark.gpu


def _run_noalloc(simd, torch, x, h, reverse, out=None):
    if out is None:
        out = torch.full((x.numel() + h.numel() - 1,), float("nan"), device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc = simd.lib.simd_b200_filter_noalloc(x.data_ptr(), x.numel(), h.data_ptr(), h.numel(),
                                           int(reverse), out.data_ptr(), C.c_void_p(stream))
    assert rc == 0
    torch.cuda.synchronize()
    return out


def test_noalloc_matches_handle_path_on_two_consecutive_calls(simd, oracle):
    import torch
    simd.use_torch_stream()
    g = torch.Generator(device="cuda").manual_seed(2024)
    # second call: shorter filter and other data, so bins left by the first call would show;
    # third: the most taps the built-in scratch holds (the transform length is forced up to
    # 16384 bins for it) on a signal that does not start on a 16-byte boundary
    cases = [(300000, 4096, False, 0), (70001, 513, True, 0), (100003, 8192, False, 3)]
    for call, (xl, hl, corr, off) in enumerate(cases):
        x = (torch.rand((xl + off,), generator=g, device="cuda") * 2 - 1)[off:]
        h = torch.rand((hl,), generator=g, device="cuda") * 2 - 1
        got = _run_noalloc(simd, torch, x, h, corr)
        with simd.Handle(xl, hl, "_overlap_save", corr) as hd:
            want = hd(x, h)
            torch.cuda.synchronize()
        xn, hn = x.cpu().numpy(), h.cpu().numpy()
        ok, r, m = parity_ok(got.cpu().numpy(), want.cpu().numpy(), xn, hn)
        assert ok, (call, r, m)
        assert rel_l2(got.cpu().numpy(), want.cpu().numpy()) <= 1e-5
        # ... and against the CPU oracle (the reference's FFT method restated), not only
        # against the library's own handle path
        truth = oracle.convolve_fft(xn, hn, reverse=1 if corr else 0)
        ok, r, m = parity_ok(got.cpu().numpy(), truth, xn, hn)
        assert ok, ("oracle", call, r, m)
    # a repeat of the last shape (kernels already loaded, output buffer already there) must
    # leave the device's free memory exactly as it found it
    out = torch.empty_like(got)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    again = _run_noalloc(simd, torch, x, h, corr, out=out)
    assert torch.cuda.mem_get_info()[0] == free_before
    assert torch.equal(again, got)


def test_noalloc_rejects_what_it_cannot_hold(simd, capfd):
    import torch
    x = torch.zeros(40000, device="cuda")
    h = torch.zeros(8193, device="cuda")
    out = torch.zeros(40000 + 8193 - 1, device="cuda")
    rc = simd.lib.simd_b200_filter_noalloc(x.data_ptr(), 40000, h.data_ptr(), 8193, 0,
                                           out.data_ptr(), None)
    assert rc == 1
    assert "exceeds" in capfd.readouterr().err
    # host memory would need a staging allocation: refused, not staged
    xh = np.zeros(100, np.float32)
    rc = simd.lib.simd_b200_filter_noalloc(xh.ctypes.data, 100, h.data_ptr(), 16, 0,
                                           out.data_ptr(), None)
    assert rc == 2
