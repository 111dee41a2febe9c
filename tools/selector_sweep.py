#!/usr/bin/env python
"""tools/selector_sweep.py -- B200 timings behind the handle-time selector (src/selector.c):
direct kernel vs fused overlap-save kernel over (batch, x, h), and the transform length N
for the overlap-save path.  Device-resident data, CUDA events, median of several launches.
Writes a markdown table to stdout (committed as profiles/selector_sweep_r01.md)."""
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import veles.simd_b200 as v  # noqa: E402


def time_call(fn, reps=7):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    return statistics.median(ts)


def run(batch, xl, hl, algorithm, force_n=None):
    x = torch.rand((batch, xl), device="cuda") * 2 - 1
    h = torch.rand((hl,), device="cuda") * 2 - 1
    y = torch.empty((batch, xl + hl - 1), device="cuda")
    if force_n:
        os.environ["VELES_B200_FORCE_N"] = str(force_n)
    hd = v.ConvolutionHandle()
    hd.algorithm, hd.x_length, hd.h_length = algorithm, xl, hl
    inner = None
    if algorithm == 2:
        inner = v.Handle(xl, hl, "_overlap_save")
        hd.handle.os = inner.handle
    os.environ.pop("VELES_B200_FORCE_N", None)
    fn = lambda: v.lib.convolve_batch(hd, x.data_ptr(), xl, h.data_ptr(), y.data_ptr(), xl + hl - 1, batch)
    us = time_call(fn)
    n = inner.transform_length if inner else 0
    if inner:
        inner.close()
    return us, n


def main():
    v.use_torch_stream()
    print("# Selector sweep (B200, device-resident, median of 7 launches, microseconds per call)\n")
    print("## direct kernel vs fused overlap-save kernel\n")
    print("| batch | x | h | direct us | overlap-save us (N) | faster |")
    print("|---|---|---|---|---|---|")
    for batch, xl in [(1, 4096), (1, 65536), (1, 1 << 20), (64, 16384), (4096, 16384)]:
        for hl in (8, 16, 32, 48, 64, 96, 128, 256, 512, 1024):
            if hl >= xl // 2:
                continue
            d, _ = run(batch, xl, hl, 0)
            o, n = run(batch, xl, hl, 2)
            print(f"| {batch} | {xl} | {hl} | {d:.1f} | {o:.1f} ({n}) | {'direct' if d <= o else 'overlap-save'} |")
    print("\n## transform length for the overlap-save path (batched, 256 MiB of input)\n")
    print("| h | " + " | ".join(f"N={n}" for n in (256, 512, 1024, 2048, 4096, 8192, 16384)) + " | model picks |")
    print("|---|" + "---|" * 8)
    for hl in (64, 128, 256, 512, 1024, 2048, 4096, 6000, 8192):
        xl = 262144
        batch = 256
        cells = []
        for n in (256, 512, 1024, 2048, 4096, 8192, 16384):
            if n < 2 * hl and not (n == 16384):
                cells.append("-")
                continue
            if n // 2 < hl and n < 16384:
                cells.append("-")
                continue
            us, _ = run(batch, xl, hl, 2, force_n=n)
            cells.append(f"{us:.0f}")
        _, pick = run(batch, xl, hl, 2)
        print(f"| {hl} | " + " | ".join(cells) + f" | {pick} |")


if __name__ == "__main__":
    main()
