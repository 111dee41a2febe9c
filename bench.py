#!/usr/bin/env python
"""bench.py -- 1D convolve / cross-correlate throughput on B200 (one JSON line on stdout).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4] [--impl b200|reference]

A "step" is one pass of the hot path over one batch of synthetic input: one call of the
batched C-ABI entry point (convolve_batch) of libSimd.so.

  value      whole-job output samples/s with x, h and result resident in HBM (device pointers
             into the same C-ABI call; CUDA-event timed, max over ranks)
  e2e        the same metric through the same call with HOST (pinned) buffers: H2D of x and
             h, the kernels, D2H of the result, all inside the timed region
  roofline   the dominant kernel's algorithmic bytes / its own CUDA-event time (measured live
             with simd_b200_profile) against the measured HBM copy bandwidth; plus the FP32
             figure, which is the co-binding roofline for these FFT/stencil kernels
  cpu_baseline  the reference's own convolve.c (oracle/_ref, + FFT stand-in) or the oracle
             port, all host cores, on a bounded sample of the same workload

Workloads (BASELINE.json configs): c4 is the headline (batched FFT/overlap-save regime, the
largest batched configuration); with no --workload the line also carries a `configs` object
with c1, c2, c3 and c5.  Under torchrun (N > 1) the headline runs the whole config on every
rank's GPU (weak scaling, no data-path collective) and configs.c4_strong the same config split
by rows over the ranks (BASELINE: "sharded by batch"; total work fixed: strong scaling), as is
c3; c5 is sharded contiguously with its (h-1)-sample halo exchanged point to point; c1/c2 run
as replicas.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (op, batch, x_len, h_len, description)
    "c1": ("convolve", 1, 1024, 50, "convolve x=1024 h=50, single signal"),
    "c2": ("correlate", 1, 65536, 512, "correlate x=65536 h=512, single signal"),
    "c3": ("convolve", 4096, 16384, 64, "batched convolve 4096 x 16384, h=64"),
    "c4": ("convolve", 1024, 262144, 4096, "batched convolve 1024 x 262144, h=4096"),
    "c5": ("convolve", 1, 2 ** 30, 8192, "single long signal x=2^30, h=8192"),
    # not a BASELINE config: the C3 batch with a 16-tap filter, where the direct kernel is HBM-bound
    "c3h16": ("convolve", 4096, 16384, 16, "batched convolve 4096 x 16384, h=16 (direct kernel, HBM-bound regime)"),
    # SURVEY 8(f) rank 1 (not a BASELINE config): extrema of a 2^28-sample series
    "peaks": ("detect_peaks", 1, 2 ** 28, 0, "detect_peaks (both kinds) on 2^28 uniform random samples"),
    "minmax": ("minmax1D", 1, 2 ** 28, 0, "minmax1D on 2^28 uniform random samples"),
    # one decimated transform level, Daubechies 8 taps, periodic extension
    "wavelet": ("wavelet_apply", 1, 2 ** 28, 8, "wavelet_apply Daubechies-8, periodic, 2^28 uniform random samples"),
}
FP32_PEAK_TFLOPS = 72.1   # measured on this pool's B200 (tools/ffma_peak, profiles/ffma_peak_r01.json)


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0}, "fallback"


def algorithmic_bytes(batch, x_len, h_len):
    return 4 * (batch * x_len + batch * (x_len + h_len - 1) + h_len)


def algorithmic_flops(alg, batch, x_len, h_len, n_fft):
    if alg == 0:
        return 2.0 * batch * x_len * h_len
    import math
    taps = min(h_len, n_fft // 2)
    parts = -(-h_len // taps)
    step = n_fft - taps + 1
    nseg = -(-(x_len + h_len - 1) // step)
    # per real segment: forward + inverse real FFT (2 * 2.5 L log2 L), spectrum product, scale
    per_seg = 2 * 2.5 * n_fft * math.log2(n_fft) + 6 * (n_fft / 2 + 1) + step
    return float(batch) * parts * nseg * per_seg


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [v.strip() for v in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        busy = [v for v in sm if mx and v > 0.5 * mx] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": mx,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_baseline(workload, sample_signals=None, target_s=2.0, threads=0):
    """Times the reference's own CPU path (oracle/_ref) -- or the oracle port when _ref is
    not present -- with all host threads on a bounded sample of the workload."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _oracle import Oracle, Reference, have_reference
    op, batch, x_len, h_len, _ = WORKLOADS[workload]
    cores = os.cpu_count() or 1
    if threads <= 0:
        threads = cores           # torchrun exports OMP_NUM_THREADS=1: ask for all cores explicitly
    if x_len > 2 ** 24:           # long single signal: time a 2^24-sample prefix (same h)
        x_len_s, nsig, note = 2 ** 24, 1, f"prefix of 2^24 samples of the signal, h={h_len}"
    else:
        x_len_s = x_len
        nsig = min(batch, sample_signals or max(1, 4 * cores))
        note = f"{nsig} of {batch} signals of {x_len} samples, h={h_len}"
    rng = np.random.default_rng(0xC0FFEE)
    x = rng.uniform(-1, 1, (nsig, x_len_s)).astype(np.float32)
    h = rng.uniform(-1, 1, h_len).astype(np.float32)
    rev = 1 if op == "correlate" else 0
    if have_reference():
        impl, kind = Reference(), "reference"
        run = lambda: impl.convolve_batch(x, h, reverse=rev, algorithm=-1, threads=threads)
    else:
        impl, kind = Oracle(), "port"
        run = lambda: impl.convolve_batch(x, h, reverse=rev, threads=threads)
    _, used = run()                       # warm-up, also tells the thread count
    outs = nsig * (x_len_s + h_len - 1)
    reps, t0 = 0, time.perf_counter()
    while True:
        run()
        reps += 1
        dt = time.perf_counter() - t0
        if dt >= target_s or reps >= 200:
            break
    used = min(used, nsig)
    return {"value": outs * reps / dt / 1e9, "unit": "GSamples/s", "cores": used, "kind": kind,
            "sample": f"{note}; {reps} repetitions in {dt:.2f} s; reference selector path "
                      f"(convolve_initialize -> convolve), one handle per thread"
                      + ("; FFT stand-in for FFTF" if kind == "reference" else "")}, outs, dt / reps


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores.
    Under torchrun only rank 0 works; other ranks exit 0."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    op, batch, x_len, h_len, desc = WORKLOADS[args.workload]
    for _ in range(max(args.warmup - 1, 0)):
        cpu_baseline(args.workload, target_s=0.2)
    times, outs = [], 0
    base = None
    for _ in range(args.steps):
        base, outs, per = cpu_baseline(args.workload, target_s=0.5)
        times.append(per)
    ms = 1e3 * statistics.mean(times)
    val = outs / (ms * 1e-3) / 1e9
    base["value"] = val
    line = {"impl": "reference", "metric": "conv1d_output_gsamples_per_s", "value": val,
            "unit": "GSamples/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: {desc}", "step": "bounded sample: " + base["sample"]},
            "cpu_baseline": base,
            "e2e": {"value": val, "unit": "GSamples/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def load_counters():
    """ncu counters of the dominant kernel of each config (DRAM bytes, warp instructions per
    launch), written by tools/ncu_counters.py from this round's captures, with the commit they
    were taken at."""
    for name in ("kernel_counters_r02.json", "traffic_r01.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            return json.load(open(p)), name
    return {}, None


def load_fp32_peak():
    p = os.path.join(ROOT, "profiles", "ffma_peak_r01.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["ffma_tflops"], "measured FFMA microbenchmark (tools/ffma_peak.cu -> profiles/ffma_peak_r01.json)"
    return 74.45, "nominal 148 SM x 128 lanes x 2 x 1.965 GHz"


class Ctx:
    pass


def bench_config(c, name, steps, warmup, min_warm_s=0.0, e2e_steps=3, do_e2e=True, do_pageable=False,
                 algorithm=-1, no_profile=False, out_row_align=1, weak=False):
    """One BASELINE config on this rank's share of it.  Batched configs are split by rows over
    the ranks (strong scaling: the batch is fixed) -- or, with weak=True, every rank runs the
    whole batch of the config on its own GPU (weak scaling: work grows with the ranks);
    single-signal configs run as replicas."""
    torch, dist, v = c.torch, c.dist, c.v
    op, batch, x_len, h_len, desc = WORKLOADS[name]
    corr = op == "correlate"
    out_len = x_len + h_len - 1
    from veles.simd_b200 import sharded
    if batch > 1 and weak and c.world > 1:
        lo, hi = 0, batch
        sharding = f"every rank runs all {batch} signals of the config on its own GPU, no collective"
    elif batch > 1:
        lo, hi = sharded.batch_rows(batch, c.rank, c.world)
        sharding = f"rows [{lo},{hi}) of {batch} on this rank, no collective" if c.world > 1 else "one GPU"
    else:
        lo, hi = 0, 1
        sharding = "replicas only (a single small signal does not shard)" if c.world > 1 else "one GPU"
    nb = hi - lo
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + c.rank)
    x = torch.rand((nb, x_len), generator=g, device="cuda") * 2 - 1
    gh = torch.Generator(device="cuda").manual_seed(0xC0FFEE)
    h = torch.rand((h_len,), generator=gh, device="cuda") * 2 - 1
    if out_row_align > 1:
        pitch = -(-out_len // out_row_align) * out_row_align
        y = torch.empty((nb, pitch), device="cuda")[:, :out_len]
    else:
        y = torch.empty((nb, out_len), device="cuda")
    if batch == 1:
        x, y = x[0], y[0]
    flush = None
    if x.numel() * 4 < 256 << 20:       # inputs smaller than L2 (126 MB): flush between steps
        flush = c.flush

    handle = v.Handle(x_len, h_len, "", corr)
    if algorithm == 0:
        os.environ["VELES_B200_NO_BATCH_CROSSOVER"] = "1"
    if algorithm >= 0 and algorithm != handle.algorithm:
        handle.close()
        handle = ForcedHandle(v, x_len, h_len, corr, algorithm)
    alg = handle.algorithm
    n_fft = handle.transform_length if alg != 0 else 0
    if alg == 1:
        n_fft = v.transform_length(x_len, h_len)

    def step():
        if flush is not None:
            flush.zero_()
        handle(x, h, out=y)

    t_warm = time.perf_counter()
    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    while time.perf_counter() - t_warm < min_warm_s:      # keep the GPU busy until clock samples exist
        step()
        torch.cuda.synchronize()
    c.barrier()
    launches0 = v.lib.simd_b200_kernel_launches()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    c.barrier()
    for e0, e1 in ev:
        if flush is not None:
            flush.zero_()
        e0.record()
        handle(x, h, out=y)
        e1.record()
    c.barrier()
    launches = v.lib.simd_b200_kernel_launches() - launches0
    # Per-kernel durations for the roofline: the same steps once more with every launch
    # bracketed by its own event pair on the same stream.  Kept out of the loop above because
    # an event record between the two launches of a step defeats their programmatic
    # dependent launch (the fused kernel starts while the spectrum kernel still runs).
    prof = {"direct": (0.0, 0), "fused": (0.0, 0), "spectrum": (0.0, 0)}
    if not no_profile:
        v.profile(True)
        v.profile_collect()
        for _ in range(steps):
            if flush is not None:
                flush.zero_()
            handle(x, h, out=y)
        torch.cuda.synchronize()
        prof = v.profile_collect()
        v.profile(False)
    step_ms = [a.elapsed_time(b) for a, b in ev]
    ms_per_step = c.max_over_ranks(sum(step_ms)) / steps
    replicas = c.world if (batch == 1 or (weak and c.world > 1)) else 1
    outputs_per_step = batch * out_len * replicas
    value = outputs_per_step / (ms_per_step * 1e-3) / 1e9

    # ---- roofline of the dominant kernel (per launch, from its own events) ----------------
    # the batched call may cross over from the handle's single-signal choice (selector.c)
    if no_profile:      # no per-kernel events: charge the whole step to the dominant kernel
        kname = "direct" if v.select_algorithm(x_len, h_len, nb) == 0 else "fused"
        prof[kname] = (sum(step_ms), steps)
    kind = "fused" if prof["fused"][1] > 0 else "direct"
    if kind == "fused" and alg == 0:
        alg = 2
        n_fft = v.transform_length(x_len, h_len)
    k_ms, k_cnt = prof[kind]
    k_avg_ms = k_ms / max(k_cnt, 1)
    launches_per_step = max(k_cnt // steps, 1)
    abytes = algorithmic_bytes(nb, x_len, h_len) / launches_per_step
    aflops = algorithmic_flops(alg, nb, x_len, h_len, n_fft) / launches_per_step
    ach_gbs = abytes / (k_avg_ms * 1e-3) / 1e9
    ach_tf = aflops / (k_avg_ms * 1e-3) / 1e12
    cnt = c.counters.get(name, {}) if c.world == 1 else {}
    traffic = cnt.get("dram_bytes_per_launch")
    fp32_peak, fp32_src = c.fp32_peak
    roofline = {"bound": "hbm", "achieved": ach_gbs, "peak": c.peaks["hbm_gbs"], "unit": "GB/s",
                "frac": ach_gbs / c.peaks["hbm_gbs"], "traffic": traffic, "peak_source": c.peak_src,
                "kernel": cnt.get("kernel", {"direct": "direct_kernel", "fused": "os_fused_kernel"}[kind]),
                "kernel_ms_per_launch": k_avg_ms, "kernel_launches_per_step": launches_per_step,
                "kernel_share_of_step": min(1.0, k_avg_ms * launches_per_step / (sum(step_ms) / steps)),
                "timing": "per-launch event pairs on the launching stream, separate pass of the same steps",
                "algorithmic_bytes_per_launch": abytes,
                "counters_from": ({"file": "profiles/" + c.counters_file, "commit": cnt.get("commit")}
                                  if traffic else None),
                "fp32": {"achieved": ach_tf, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": ach_tf / fp32_peak, "peak_source": fp32_src,
                         "algorithmic_flops_per_launch": aflops}}
    if cnt.get("warp_instructions_per_launch"):
        # instruction-issue roofline: the kernel's own warp instructions at one per clock per
        # scheduler (4 per SM); the FP pipe sustains 3.88 of those 4 (tools/fp_pipes.cu)
        sm_clk = 148 * 4 * (c.peaks.get("sm_max_mhz", 1965.0) * 1e6)
        floor_ms = cnt["warp_instructions_per_launch"] / sm_clk * 1e3
        roofline["issue_frac"] = floor_ms / k_avg_ms
        roofline["issue"] = {"warp_instructions_per_launch": cnt["warp_instructions_per_launch"],
                             "floor_ms": floor_ms, "note": "time at one warp instruction per clock per scheduler / measured kernel time"}
    if batch == 1 and x_len < 1 << 20:
        roofline["note"] = "single small signal: launch/latency-bound, the fraction is not meaningful; see us_per_call"

    # ---- end to end: same C-ABI call, HOST buffers ------------------------------------------
    e2e = None
    if do_e2e:
        shape_x = (nb, x_len) if batch > 1 else (x_len,)
        shape_y = (nb, out_len) if batch > 1 else (out_len,)
        xh = torch.empty(shape_x, dtype=torch.float32).pin_memory()
        hh = torch.empty((h_len,), dtype=torch.float32).pin_memory()
        yh = torch.empty(shape_y, dtype=torch.float32).pin_memory()
        if x.numel() < 2 ** 28:
            xh.copy_(x.cpu())
        else:
            xh.uniform_(-1, 1)
        hh.copy_(h.cpu())
        reps = e2e_steps if x.numel() >= 1 << 22 else 200          # tiny calls: time many
        handle(xh, hh, out=yh)            # warm-up (also sizes the staging buffers)
        c.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            handle(xh, hh, out=yh)        # returns with the result in host memory
        dt = c.max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": outputs_per_step * reps / dt / 1e9, "unit": "GSamples/s",
               "h2d_bytes_per_step": 4 * (nb * x_len + h_len) * c.world,
               "d2h_bytes_per_step": 4 * nb * out_len * c.world,
               "steps": reps, "ms_per_step": 1e3 * dt / reps, "host_memory": "pinned",
               "result_checksum": float(yh.reshape(-1)[:1024].double().sum())}
        if batch == 1 and x_len < 1 << 20:
            e2e["us_per_call"] = 1e6 * dt / reps
        if do_pageable:
            # what a caller of the reference API hands over: plain malloc'ed memory
            import numpy as np
            xp = np.empty(shape_x, np.float32)
            xp[...] = xh.numpy()
            yp = np.empty(shape_y, np.float32)
            hp = hh.numpy().copy()
            handle(xp, hp, out=yp)
            c.barrier()
            t0 = time.perf_counter()
            for _ in range(max(reps // 2, 2) if reps < 50 else reps):
                handle(xp, hp, out=yp)
            n_p = max(reps // 2, 2) if reps < 50 else reps
            dtp = c.max_over_ranks(time.perf_counter() - t0)
            assert float(abs(yp.reshape(-1)[:1024].astype("float64").sum() - e2e["result_checksum"])) < 1e-3
            e2e["pageable"] = {"value": outputs_per_step * n_p / dtp / 1e9, "unit": "GSamples/s",
                               "ms_per_step": 1e3 * dtp / n_p, "steps": n_p,
                               "host_memory": "pageable (numpy / malloc), what a reference caller passes"}
            del xp, yp
        del xh, yh
    handle.close()
    res = {"value": value, "unit": "GSamples/s", "ms_per_step": ms_per_step,
           "workload": f"{name}: {desc}", "op": op, "signals_total": batch, "signals_this_rank": nb,
           "x_length": x_len, "h_length": h_len,
           "algorithm": {0: "direct", 1: "fft", 2: "overlap_save"}[alg], "transform_length": n_fft,
           "scaling": ("weak" if weak else "strong") if batch > 1 else "replicas",
           "sharding": sharding,
           "l2": ("inputs larger than L2" if flush is None else "256 MiB flush buffer written between steps"),
           "gpu_launches": int(launches), "roofline": roofline, "e2e": e2e,
           "step_ms": {"min": min(step_ms), "max": max(step_ms)}}
    if batch == 1 and x_len < 1 << 20:
        res["us_per_call"] = ms_per_step * 1e3
    del x, y
    torch.cuda.empty_cache()
    return res


def bench_c5_sharded(c, steps, warmup):
    """C5 on N GPUs: the 2^30-sample signal is pre-sharded (contiguous, device resident);
    each step = the (h-1)-sample halo exchange (NCCL send/recv) + the shard kernels.
    Total work is fixed: strong scaling."""
    torch, dist, v = c.torch, c.dist, c.v
    from veles.simd_b200 import sharded
    op, batch, x_len, h_len, desc = WORKLOADS["c5"]
    s0, e0 = sharded.signal_range(x_len, c.rank, c.world)
    ext, x_local = sharded.alloc_shard(e0 - s0, h_len - 1, device="cuda")
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + c.rank)
    x_local.copy_(torch.rand((e0 - s0,), generator=g, device="cuda") * 2 - 1)
    gh = torch.Generator(device="cuda").manual_seed(0xC0FFEE)
    h = torch.rand((h_len,), generator=gh, device="cuda") * 2 - 1
    n_out = (e0 - s0) + (h_len - 1 if c.rank == c.world - 1 else 0)
    y = torch.empty((n_out,), device="cuda")

    def step():
        sharded.filter_long_signal(x_local, h, x_len, c.rank, c.world, sharded.gpu_compute, ext=ext, out=y,
                                   overlap=bool(os.environ.get("VELES_B200_OVERLAP_HALO")))

    for _ in range(warmup):
        step()
    c.barrier()
    launches0 = v.lib.simd_b200_kernel_launches()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        step()
    ev1.record()
    c.barrier()
    ms = c.max_over_ranks(ev0.elapsed_time(ev1)) / steps
    # the exchange alone (same loop without the kernels), to say what it costs per step
    xe0, xe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c.barrier()
    xe0.record()
    for _ in range(steps):
        sharded.exchange_halo(x_local, h_len - 1, c.rank, c.world, into=ext[:h_len - 1] if c.rank > 0 else None)
    xe1.record()
    c.barrier()
    ex_ms = c.max_over_ranks(xe0.elapsed_time(xe1)) / steps
    res = {"value": (x_len + h_len - 1) / (ms * 1e-3) / 1e9, "unit": "GSamples/s", "ms_per_step": ms,
           "workload": f"c5: {desc}", "scaling": "strong",
           "sharding": "contiguous shards + (h-1)-sample halo via NCCL send/recv, pre-sharded device-resident data",
           "halo_bytes_per_rank": 4 * (h_len - 1), "halo_exchange_ms": ex_ms,
           "l2": "inputs larger than L2",
           "gpu_launches": int(v.lib.simd_b200_kernel_launches() - launches0), "e2e": None, "roofline": None}
    del x_local, ext, y
    torch.cuda.empty_cache()
    return res


def e2e_fewer_bytes(c, steps=2):
    """The two PCIe-saving chains (simd/b200.h), end to end with pinned host buffers, next to the
    plain calls they replace: (1) C4 with int16 samples widened on the device; (2) correlate a
    2^26-sample signal with a 512-tap template and return only the extrema."""
    import numpy as np
    torch, v = c.torch, c.v
    out = {}
    # (1) int16 input, C4 shape
    _, batch, x_len, h_len, _ = WORKLOADS["c4"]
    out_len = x_len + h_len - 1
    x16 = torch.empty((batch, x_len), dtype=torch.int16).pin_memory()
    x16.random_(-32768, 32767)
    hh = (torch.rand((h_len,)) * 2 - 1).pin_memory()
    yh = torch.empty((batch, out_len), dtype=torch.float32).pin_memory()
    with v.Handle(x_len, h_len) as hd:
        v.batch_int16(hd, x16, hh, out=yh)
        t0 = time.perf_counter()
        for _ in range(steps):
            v.batch_int16(hd, x16, hh, out=yh)
        dt = (time.perf_counter() - t0) / steps
    out["c4_int16_in"] = {"value": batch * out_len / dt / 1e9, "unit": "GSamples/s", "ms_per_step": 1e3 * dt,
                          "h2d_bytes_per_step": 2 * batch * x_len + 4 * h_len, "d2h_bytes_per_step": 4 * batch * out_len,
                          "pcie_bytes_per_output": (2 * batch * x_len + 4 * batch * out_len) / (batch * out_len),
                          "call": "convolve_batch_int16 (int16 samples cross PCIe, widened on the device)",
                          "replaces": "int16_to_float on the host + convolve_batch: 8 PCIe bytes per output"}
    del x16, yh
    # (2) correlate -> peaks
    xl, hl = 2 ** 26, 512
    xs = torch.empty((xl,), dtype=torch.float32).pin_memory()
    t = torch.arange(xl, dtype=torch.float32)
    xs.copy_(torch.sin(t * 0.001) + 0.25 * torch.sin(t * 0.0137) + 0.01 * torch.randn(xl))
    del t
    hs = torch.hann_window(hl).pin_memory()
    pos, val = v.cross_correlate_peaks(xs, hs)
    t0 = time.perf_counter()
    for _ in range(steps):
        pos, val = v.cross_correlate_peaks(xs, hs)
    dt = (time.perf_counter() - t0) / steps
    ys = torch.empty((xl + hl - 1,), dtype=torch.float32).pin_memory()
    with v.Handle(xl, hl, "", True) as hd:
        hd(xs, hs, out=ys)
        t0 = time.perf_counter()
        for _ in range(steps):
            hd(xs, hs, out=ys)
            pos2, val2 = v.detect_peaks(ys.numpy())
        dt2 = (time.perf_counter() - t0) / steps
    assert np.array_equal(pos, pos2) and np.array_equal(val, val2)
    out["correlate_peaks_only"] = {
        "value": (xl + hl - 1) / dt / 1e9, "unit": "GSamples/s (correlation samples examined)", "ms_per_step": 1e3 * dt,
        "h2d_bytes_per_step": 4 * (xl + hl), "d2h_bytes_per_step": 8 * int(pos.size), "extrema": int(pos.size),
        "pcie_bytes_per_output": (4 * (xl + hl) + 8 * int(pos.size)) / (xl + hl - 1),
        "call": "simd_b200_cross_correlate_peaks (the correlation stays in HBM, {position, value} pairs come back)",
        "two_call_chain": {"value": (xl + hl - 1) / dt2 / 1e9, "ms_per_step": 1e3 * dt2,
                           "what": "cross_correlate to host memory, then detect_peaks on that host array "
                                   "(result down, result up again, peaks down): same extrema"}}
    return out


def e2e_single_process(c, name, steps):
    """Rank 0 alone: the whole batch in host memory through one convolve_batch call that spreads
    it over all the GPUs of the box (simd_b200_use_devices), other ranks idle."""
    torch, v = c.torch, c.v
    op, batch, x_len, h_len, desc = WORKLOADS[name]
    out_len = x_len + h_len - 1
    import numpy as np
    # the library's own pinned allocator (portable: every GPU's copy engine reads it in place)
    xa, ya = v.PinnedArray((batch, x_len)), v.PinnedArray((batch, out_len))
    xh, yh = xa.array, ya.array
    rng = np.random.default_rng(0xC0FFEE)
    for b0 in range(0, batch, 64):
        xh[b0:b0 + 64] = rng.uniform(-1, 1, xh[b0:b0 + 64].shape).astype(np.float32)
    hh = rng.uniform(-1, 1, h_len).astype(np.float32)
    n = v.use_devices(c.world)
    with v.Handle(x_len, h_len, "", op == "correlate") as hd:
        hd(xh, hh, out=yh)
        t0 = time.perf_counter()
        for _ in range(steps):
            hd(xh, hh, out=yh)
        dt = time.perf_counter() - t0
    v.use_devices(1)
    xa.close(); ya.close()
    return {"value": batch * out_len * steps / dt / 1e9, "unit": "GSamples/s", "ms_per_step": 1e3 * dt / steps,
            "devices": n, "steps": steps, "h2d_bytes_per_step": 4 * (batch * x_len + h_len),
            "d2h_bytes_per_step": 4 * batch * out_len, "host_memory": "pinned",
            "how": "one process, one convolve_batch call, simd_b200_use_devices(N): one worker thread + pipeline per GPU"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS),
                    help="one workload only (default: c4 as the headline plus every other BASELINE config)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="headline only")
    ap.add_argument("--strong", action="store_true",
                    help="N > 1: split the headline config's batch by rows instead of one whole batch per rank")
    ap.add_argument("--out-row-align", type=int, default=1,
                    help="pad result rows to a multiple of this many floats (experiment)")
    ap.add_argument("--no-profile", action="store_true",
                    help="do not bracket kernels with events (they sit between the two launches of a step)")
    ap.add_argument("--min-warm-s", type=float, default=0.6,
                    help="keep stepping this long before the timed region so clock samples exist")
    ap.add_argument("--algorithm", type=int, default=-1,
                    help="force 0 direct / 2 overlap-save (profiling only)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    all_configs = args.workload is None and not args.no_configs
    if args.workload is None:
        args.workload = "c4"
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    import veles.simd_b200 as v

    c = Ctx()
    c.torch, c.dist, c.v = torch, dist, v
    c.world = int(os.environ.get("WORLD_SIZE", "1"))
    c.rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    if c.world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    v.use_torch_stream()
    c.peaks, c.peak_src = load_peaks()
    c.counters, c.counters_file = load_counters()
    c.fp32_peak = load_fp32_peak()
    c.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if c.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(val):
        if c.world == 1:
            return val
        t = torch.tensor([val], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    c.barrier, c.max_over_ranks = barrier, max_over_ranks
    world, rank = c.world, c.rank

    if args.workload == "peaks":
        return run_detect_peaks(args, v, world, rank)
    if args.workload == "minmax":
        return run_minmax(args, v, world, rank)
    if args.workload == "wavelet":
        return run_wavelet(args, v, world, rank)

    # clocks are sampled from before the warm-up to the end of the timed region (the timed
    # region alone can be shorter than one nvidia-smi sampling period)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    if args.workload == "c5" and world > 1:
        head = bench_c5_sharded(c, args.steps, args.warmup)
    else:
        # headline at N > 1: every rank the whole config on its own GPU (weak scaling, as in round
        # 1); the same config split by rows (strong scaling, BASELINE's "sharded by batch") is in
        # configs.<name>_strong right beside it
        head = bench_config(c, args.workload, args.steps, args.warmup, min_warm_s=args.min_warm_s,
                            e2e_steps=args.e2e_steps, do_e2e=not args.no_e2e,
                            do_pageable=not args.no_e2e and world == 1,
                            algorithm=args.algorithm, no_profile=args.no_profile,
                            out_row_align=args.out_row_align, weak=not args.strong)
    clocks = sampler.stop() if rank == 0 else None

    configs = {}
    if world > 1 and WORKLOADS[args.workload][1] > 1 and not args.strong and not args.no_configs:
        configs[args.workload + "_strong"] = bench_config(c, args.workload, args.steps, args.warmup,
                                                          e2e_steps=2, do_e2e=not args.no_e2e)
    if all_configs:
        # every other BASELINE config, fewer steps each; at N > 1 the batched ones are split by
        # rows (strong scaling), C5 is sharded with its halo, C1/C2 run as replicas
        for name in ("c1", "c2", "c3", "c5"):
            st = max(10, min(args.steps, 30)) if name != "c5" else max(5, min(args.steps, 10))
            try:
                if name == "c5" and world > 1:
                    configs[name] = bench_c5_sharded(c, st, args.warmup)
                else:
                    configs[name] = bench_config(c, name, st, args.warmup, e2e_steps=2, do_e2e=not args.no_e2e)
            except Exception as e:          # one config must not take the headline down
                configs[name] = {"error": f"{type(e).__name__}: {e}"}
    if all_configs and world == 1 and not args.no_e2e:
        try:
            configs.update(e2e_fewer_bytes(c))
        except Exception as e:
            configs["fewer_bytes"] = {"error": f"{type(e).__name__}: {e}"}
    single = None
    if world > 1 and not args.no_e2e and WORKLOADS[args.workload][1] > 1:
        # the other ranks must leave their GPUs idle while rank 0 drives all of them: wait on a
        # CPU (gloo) barrier -- an NCCL barrier is a kernel spinning on every waiting GPU
        cpu_group = dist.new_group(backend="gloo")
        barrier()
        if rank == 0:
            try:
                single = e2e_single_process(c, args.workload, args.e2e_steps)
            except Exception as e:
                single = {"error": f"{type(e).__name__}: {e}"}
        dist.barrier(group=cpu_group)

    base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        base, _, _ = cpu_baseline(args.workload)
        base["fft_note"] = fft_standin_note()
        if all_configs:
            for name in configs:
                if name in WORKLOADS and "error" not in configs[name]:
                    b, _, _ = cpu_baseline(name, target_s=1.0)
                    configs[name]["cpu_baseline"] = b

    if rank == 0:
        line = {"metric": "conv1d_output_gsamples_per_s", "value": head["value"], "unit": "GSamples/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": head["ms_per_step"], "higher_is_better": True,
                "scaling": head["scaling"] if head["scaling"] != "replicas" else "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {k: head[k] for k in ("workload", "op", "signals_total", "signals_this_rank", "x_length",
                                                "h_length", "algorithm", "transform_length", "l2", "sharding")
                           if k in head},
                "clocks": clocks, "e2e": head.get("e2e"), "gpu_launches": head["gpu_launches"],
                "roofline": head.get("roofline"), "cpu_baseline": base}
        for k in ("step_ms", "halo_bytes_per_rank", "halo_exchange_ms", "us_per_call"):
            if k in head:
                line[k] = head[k]
        if single is not None:
            line["e2e_single_process"] = single
        if configs:
            line["configs"] = configs
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def fft_standin_note():
    """How fast the CPU arm's FFT is, so that a reader can scale the ratio to a SIMD FFT:
    single-core MFLOP/s (5 N log2 N convention) of the stand-in on a 16384-point real transform
    pair, via the oracle's overlap-save call."""
    try:
        import numpy as np
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from _oracle import Reference, have_reference, Oracle
        impl = Reference() if have_reference() else Oracle()
        x = np.random.default_rng(0).uniform(-1, 1, 1 << 20).astype(np.float32)
        h = np.random.default_rng(1).uniform(-1, 1, 4096).astype(np.float32)
        run = (lambda: impl.convolve_batch(x[None, :], h, reverse=0, algorithm=-1, threads=1)) if have_reference() \
            else (lambda: impl.convolve_batch(x[None, :], h, reverse=0, threads=1))
        run()
        t0 = time.perf_counter()
        run()
        dt = time.perf_counter() - t0
        L, step = 16384, 16384 - 4095
        nseg = -(-(x.size + h.size - 1) // step)
        flops = nseg * 2 * 2.5 * L * 14
        return {"single_core_mflops": flops / dt / 1e6,
                "what": "FFTF is absent from the reference tree; the CPU arm transforms with oracle/fft_pow2.h, "
                        "scalar radix-4 C at -O2. A SIMD FFT library typically sustains several GFLOP/s per "
                        "core: scale the GPU/CPU ratios down by that factor over single_core_mflops."}
    except Exception as e:      # the note must never break the bench line
        return {"error": str(e)}


def run_detect_peaks(args, v, world, rank):
    """detect_peaks, device-resident: value = input samples/s; roofline on the algorithmic
    bytes 4*size (series read once) + 8*peaks (list written once) of the whole three-kernel
    call, timed with CUDA events on the launching stream."""
    import numpy as np
    import torch
    _, _, size, _, desc = WORKLOADS["peaks"]
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + rank)
    data = torch.rand((size,), generator=g, device="cuda") * 2 - 1
    out = torch.empty((size, 2), dtype=torch.int32, device="cuda")
    for _ in range(args.warmup):
        n = v.detect_peaks_device(data, out, 3)
    torch.cuda.synchronize()
    launches0 = v.lib.simd_b200_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        n = v.detect_peaks_device(data, out, 3)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    peaks, _ = load_peaks()
    abytes = 4.0 * size + 8.0 * n
    ach = abytes / (ms * 1e-3) / 1e9
    base = None
    if not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from _oracle import Oracle, Reference, have_reference
        sample = (np.random.default_rng(1).uniform(-1, 1, 2 ** 24)).astype(np.float32)
        impl, kind = (Reference(), "reference") if have_reference() else (Oracle(), "port")
        t0, reps = time.perf_counter(), 0
        while time.perf_counter() - t0 < 2.0:
            impl.lib.detect_peaks if kind == "reference" else None
            (impl.detect_peaks(sample[: 2 ** 20], 3, 1) if kind == "reference" else impl.detect_peaks(sample[: 2 ** 20], 3))
            reps += 1
        dt = time.perf_counter() - t0
        base = {"value": reps * 2 ** 20 / dt / 1e9, "unit": "GSamples/s", "cores": 1, "kind": kind,
                "sample": f"2^20-sample prefix, {reps} repetitions in {dt:.2f} s, single thread "
                          "(the reference is single-threaded; includes the ctypes result copy)"}
    line = {"metric": "detect_peaks_input_gsamples_per_s", "value": size / (ms * 1e-3) / 1e9,
            "unit": "GSamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"peaks: {desc}", "extrema_found": int(n), "l2": "inputs larger than L2"},
            "gpu_launches": int(v.lib.simd_b200_kernel_launches() - launches0), "e2e": None,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": ach / peaks["hbm_gbs"], "traffic": None,
                         "kernel": "peaks_count + peaks_sums + peaks_write (whole call)",
                         "algorithmic_bytes_per_launch": abytes},
            "cpu_baseline": base, "clocks": None}
    if rank == 0:
        print(json.dumps(line), flush=True)


def run_minmax(args, v, world, rank):
    """minmax1D, device-resident: algorithmic bytes = 4 per sample (read once)."""
    import numpy as np
    import torch
    _, _, size, _, desc = WORKLOADS["minmax"]
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + rank)
    data = torch.rand((size,), generator=g, device="cuda") * 2 - 1
    out = torch.zeros(2, device="cuda")
    call = lambda: v.lib.simd_b200_minmax1D_device(data.data_ptr(), size, out.data_ptr())
    for _ in range(args.warmup):
        call()
    torch.cuda.synchronize()
    launches0 = v.lib.simd_b200_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    assert tuple(out.cpu().tolist()) == (float(data.min()), float(data.max()))
    peaks, _ = load_peaks()
    ach = 4.0 * size / (ms * 1e-3) / 1e9
    base = None
    if not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from _oracle import Oracle, Reference, have_reference
        sample = np.random.default_rng(1).uniform(-1, 1, 2 ** 24).astype(np.float32)
        impl, kind = (Reference(), "reference") if have_reference() else (Oracle(), "port")
        t0, reps = time.perf_counter(), 0
        while time.perf_counter() - t0 < 2.0:
            impl.minmax1D(sample)
            reps += 1
        dt = time.perf_counter() - t0
        base = {"value": reps * sample.size / dt / 1e9, "unit": "GSamples/s", "cores": 1, "kind": kind,
                "sample": f"2^24 samples, {reps} repetitions in {dt:.2f} s, single thread (AVX path)"}
    line = {"metric": "minmax1d_input_gsamples_per_s", "value": size / (ms * 1e-3) / 1e9,
            "unit": "GSamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": f"minmax: {desc}",
                                                             "l2": "inputs larger than L2"},
            "gpu_launches": int(v.lib.simd_b200_kernel_launches() - launches0), "e2e": None,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": ach / peaks["hbm_gbs"], "traffic": None,
                         "kernel": "minmax_partial_kernel (+ one-CTA final)",
                         "algorithmic_bytes_per_launch": 4.0 * size},
            "cpu_baseline": base, "clocks": None}
    if rank == 0:
        print(json.dumps(line), flush=True)


def run_wavelet(args, v, world, rank):
    """wavelet_apply, device-resident: value = input samples/s; algorithmic bytes = 8 per
    input sample (read once, two half-length outputs written once)."""
    import numpy as np
    import torch
    _, _, size, order, desc = WORKLOADS["wavelet"]
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + rank)
    data = torch.rand((size,), generator=g, device="cuda") * 2 - 1
    hi = torch.empty(size // 2, device="cuda")
    lo = torch.empty(size // 2, device="cuda")
    call = lambda: v.lib.wavelet_apply(v.WAVELET_DAUBECHIES, order, v.EXTENSION_PERIODIC, data.data_ptr(),
                                       size, hi.data_ptr(), lo.data_ptr())
    for _ in range(args.warmup):
        call()
    torch.cuda.synchronize()
    launches0 = v.lib.simd_b200_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    launches = v.lib.simd_b200_kernel_launches() - launches0
    # orthonormal pair + periodic extension: the transform keeps the energy
    e_in = float((data.double() ** 2).sum())
    e_out = float((hi.double() ** 2).sum() + (lo.double() ** 2).sum())
    assert abs(e_out / e_in - 1) < 1e-5, (e_in, e_out)
    # end to end through the reference-facing call with host buffers (pinned): H2D of the
    # series, kernel, D2H of both halves inside the timed region (one stream, not pipelined)
    e2e = None
    if not args.no_e2e:
        hx = torch.empty(size, dtype=torch.float32, pin_memory=True)
        hx.copy_(data)
        hhi = torch.empty(size // 2, dtype=torch.float32, pin_memory=True)
        hlo = torch.empty(size // 2, dtype=torch.float32, pin_memory=True)
        hcall = lambda: v.lib.wavelet_apply(v.WAVELET_DAUBECHIES, order, v.EXTENSION_PERIODIC, hx.data_ptr(),
                                            size, hhi.data_ptr(), hlo.data_ptr())
        hcall()
        t0 = time.perf_counter()
        for _ in range(3):
            hcall()                      # returns when the results are in host memory
        dt = (time.perf_counter() - t0) / 3
        assert torch.equal(hhi, hi.cpu()) and torch.equal(hlo, lo.cpu())
        e2e = {"value": size / dt / 1e9, "unit": "GSamples/s", "h2d_bytes_per_step": 4 * size,
               "d2h_bytes_per_step": 4 * size, "steps": 3}
        del hx, hhi, hlo
    peaks, _ = load_peaks()
    ach = 8.0 * size / (ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic_r01.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get("wavelet", {}).get("dram_bytes_per_launch")
    base = None
    if not args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from _oracle import Oracle, Reference, have_reference
        sample = np.random.default_rng(1).uniform(-1, 1, 2 ** 22).astype(np.float32)
        if have_reference():
            ref, kind = Reference(), "reference"
            run = lambda: ref.wavelet_apply(0, order, 0, sample, na=True)
            how = "wavelet_apply_na (scalar loop; the AVX entry point needs its prepared layout)"
        else:
            orc, kind = Oracle(), "port"
            row = v.wavelet_filters(v.WAVELET_DAUBECHIES, order)[1]
            run = lambda: orc.wavelet_apply(row, 0, sample)
            how = "C restatement of wavelet_apply_na"
        t0, reps = time.perf_counter(), 0
        while time.perf_counter() - t0 < 2.0:
            run()
            reps += 1
        dt = time.perf_counter() - t0
        base = {"value": reps * sample.size / dt / 1e9, "unit": "GSamples/s", "cores": 1, "kind": kind,
                "sample": f"2^22 samples, {reps} repetitions in {dt:.2f} s, single thread, {how}"}
    line = {"metric": "wavelet_input_gsamples_per_s", "value": size / (ms * 1e-3) / 1e9,
            "unit": "GSamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": f"wavelet: {desc}",
                                                             "l2": "inputs larger than L2"},
            "gpu_launches": int(launches), "e2e": e2e,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": ach / peaks["hbm_gbs"], "traffic": traffic,
                         "kernel": "wavelet_decimated_kernel",
                         "note": "the peak is the measured copy bandwidth; this kernel runs slightly above it",
                         "algorithmic_bytes_per_launch": 8.0 * size},
            "cpu_baseline": base, "clocks": None}
    if rank == 0:
        print(json.dumps(line), flush=True)


class ForcedHandle:
    """Profiling aid: run a workload with a chosen algorithm through the explicit handle
    families (convolve_overlap_save_* / *_simd) instead of the selector."""

    def __init__(self, v, x_len, h_len, corr, algorithm):
        self.v, self.algorithm, self.corr = v, algorithm, corr
        self.x_len, self.h_len = x_len, h_len
        self.inner = v.Handle(x_len, h_len, "_overlap_save", corr) if algorithm == 2 else None
        self.transform_length = self.inner.transform_length if self.inner else None

    def __call__(self, x, h, out=None):
        # batch through the best-method struct carrying the forced algorithm
        v = self.v
        hd = v.ConvolutionHandle()
        hd.algorithm, hd.x_length, hd.h_length = self.algorithm, self.x_len, self.h_len
        if self.inner:
            hd.handle.os = self.inner.handle
        fn = v.lib.cross_correlate_batch if self.corr else v.lib.convolve_batch
        xs = x.stride(0) if x.dim() == 2 else x.shape[-1]
        os_ = out.stride(0) if out.dim() == 2 else out.shape[-1]
        fn(hd, x.data_ptr(), xs, h.data_ptr(), out.data_ptr(), os_, x.shape[0] if x.dim() == 2 else 1)
        return out

    def close(self):
        if self.inner:
            self.inner.close()


if __name__ == "__main__":
    main()
