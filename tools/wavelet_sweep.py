#!/usr/bin/env python
"""tools/wavelet_sweep.py -- device-resident throughput of the wavelet kernels over filter
length, family and (stationary) level; prints a markdown table.  2^27 input samples, CUDA
events on the library stream, 20 timed calls after 3 warm-ups."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import veles.simd_b200 as v  # noqa: E402

v.use_torch_stream()
n = 1 << 27
x = torch.rand(n, device="cuda") * 2 - 1
hi, lo = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
peak = 6550.4
print("| transform | taps | level | ms | input GS/s | GB/s (algorithmic) | of HBM peak | TFLOP/s |")
print("|---|---|---|---|---|---|---|---|")
cases = [(0, o, None) for o in (2, 4, 8, 12, 16, 20, 38, 76)] + [(2, 16, None), (1, 30, None)] + \
        [(0, 8, l) for l in (1, 2, 4, 8)] + [(0, 76, 1), (0, 76, 10)]
for type_, order, level in cases:
    if level is None:
        call = lambda: v.lib.wavelet_apply(type_, order, 0, x.data_ptr(), n, hi.data_ptr(), lo.data_ptr())
        nbytes, flops = 8.0 * n, 2.0 * order * n
    else:
        call = lambda: v.lib.stationary_wavelet_apply(type_, order, level, 0, x.data_ptr(), n, hi.data_ptr(),
                                                      lo.data_ptr())
        nbytes, flops = 12.0 * n, 4.0 * order * n
    for _ in range(3):
        call()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    name = ("decimated " if level is None else "stationary ") + ["Daubechies", "Coiflet", "Symlet"][type_]
    print(f"| {name} | {order} | {level or '-'} | {ms:.3f} | {n / ms / 1e6:.1f} | {nbytes / ms / 1e6:.0f} | "
          f"{nbytes / ms / 1e6 / peak:.2f} | {flops / ms / 1e9:.1f} |")
