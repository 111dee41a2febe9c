#!/usr/bin/env python
"""tools/selector_sweep.py -- B200 timings behind the handle-time selector (src/selector.c):
direct kernel vs fused overlap-save kernel over (batch, x, h), and the transform length N
for the overlap-save path.  Device-resident data, CUDA events, median of several launches.
Writes a markdown table to stdout (committed as profiles/selector_sweep_r02.md)."""
import os
import statistics
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import veles.simd_b200 as v  # noqa: E402


def time_call(fn, reps=7):
    """(call_us, kernels_us): median over `reps` of the time between two events around one call
    (includes the launch latency of a cold stream, which is all a tiny job costs) and of the
    summed device time of the call's kernels, each bracketed by its own event pair
    (simd_b200_profile) -- the figure that says what the kernels themselves cost."""
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts, ks = [], []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    v.profile(True)
    v.profile_collect()
    for _ in range(reps):
        fn()
        torch.cuda.synchronize()
        prof = v.profile_collect()
        ks.append(1e3 * sum(ms for ms, _ in prof.values()))
    v.profile(False)
    return statistics.median(ts), statistics.median(ks)


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
    elif algorithm == 1:
        inner = v.Handle(xl, hl, "_fft")
        hd.handle.fft = inner.handle
    os.environ.pop("VELES_B200_FORCE_N", None)
    if algorithm == 0:
        os.environ["VELES_B200_NO_BATCH_CROSSOVER"] = "1"
    fn = lambda: v.lib.convolve_batch(hd, x.data_ptr(), xl, h.data_ptr(), y.data_ptr(), xl + hl - 1, batch)
    us, kus = time_call(fn)
    os.environ.pop("VELES_B200_NO_BATCH_CROSSOVER", None)
    n = inner.transform_length if inner else 0
    if inner:
        inner.close()
    return (us, kus), n


def main():
    v.use_torch_stream()
    print("# Selector sweep (B200, device-resident, median of 7 launches, microseconds)\n")
    print("Two figures per cell: `call / kernels` = time between two events around one call (a cold stream's "
          "launch latency included: what a single call costs) / summed device time of the call's kernels, each "
          "between its own event pair (what the kernels cost).\n")
    print("## the reference's own benchmark shapes (tests/convolve.cc:168-400)\n")
    print("These are the shapes whose CPU timings produced the reference's 200 / 350 thresholds "
          "(src/convolve.c:349-363).  `ref picks` = its rule, `B200 picks` = src/selector.c.\n")
    print("| x | h | direct | FFT (N) | overlap-save (N) | ref picks | B200 picks |")
    print("|---|---|---|---|---|---|---|")
    names = {0: "brute", 1: "FFT", 2: "overlap-save"}
    for xl, hl in [(512, 512), (256, 256), (128, 128), (350, 350), (351, 350), (1000, 50), (2000, 950), (200, 50),
                   (201, 50), (1024, 50)]:
        d, _ = run(1, xl, hl, 0)
        f, nf = run(1, xl, hl, 1)
        cell_o = "-"
        if hl < xl // 2:
            o, no = run(1, xl, hl, 2)
            cell_o = f"{o[0]:.1f} / {o[1]:.1f} ({no})"
        ref = (2 if xl > 200 else 0) if xl > 2 * hl else (1 if xl > 350 else 0)
        print(f"| {xl} | {hl} | {d[0]:.1f} / {d[1]:.1f} | {f[0]:.1f} / {f[1]:.1f} ({nf}) | {cell_o} | {names[ref]} | "
              f"{names[v.select_algorithm(xl, hl)]} |")
    print("\n## direct kernel vs fused overlap-save kernel\n")
    print("| batch | x | h | direct call / kernels | overlap-save call / kernels (N) | faster (kernels) |")
    print("|---|---|---|---|---|---|")
    for batch, xl in [(1, 4096), (1, 65536), (1, 1 << 20), (64, 16384), (4096, 16384)]:
        for hl in (8, 16, 32, 48, 64, 96, 128, 256, 512, 1024):
            if hl >= xl // 2:
                continue
            d, _ = run(batch, xl, hl, 0)
            o, n = run(batch, xl, hl, 2)
            print(f"| {batch} | {xl} | {hl} | {d[0]:.1f} / {d[1]:.1f} | {o[0]:.1f} / {o[1]:.1f} ({n}) | "
                  f"{'direct' if d[1] <= o[1] else 'overlap-save'} |")
    print("\n## transform length for the overlap-save path (batched, 256 MiB of input; kernel time)\n")
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
            cells.append(f"{us[1]:.0f}")
        _, pick = run(batch, xl, hl, 2)
        print(f"| {hl} | " + " | ".join(cells) + f" | {pick} |")


if __name__ == "__main__":
    main()
