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

Workloads (BASELINE.json configs): c4 is the default headline (batched FFT/overlap-save
regime, the largest batched configuration); c1/c2/c3/c5 are selectable for profiling.
Under torchrun (N > 1) every rank processes the full per-GPU workload on its own GPU
(independent signals, no data-path collective): weak scaling.
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import torch.distributed as dist
    import veles.simd_b200 as v

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    v.use_torch_stream()

    op, batch, x_len, h_len, desc = WORKLOADS[args.workload]
    corr = op == "correlate"
    out_len = x_len + h_len - 1
    if args.workload == "c5" and world > 1:
        return run_sharded_long_signal(args, v, world, rank, local)
    if args.workload == "peaks":
        return run_detect_peaks(args, v, world, rank)
    if args.workload == "minmax":
        return run_minmax(args, v, world, rank)
    if args.workload == "wavelet":
        return run_wavelet(args, v, world, rank)
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + rank)
    x = torch.rand((batch, x_len), generator=g, device="cuda") * 2 - 1
    h = torch.rand((h_len,), generator=g, device="cuda") * 2 - 1
    if args.out_row_align > 1:
        pitch = -(-out_len // args.out_row_align) * args.out_row_align
        y = torch.empty((batch, pitch), device="cuda")[:, :out_len]
    else:
        y = torch.empty((batch, out_len), device="cuda")
    flush = None
    in_bytes = x.numel() * 4
    if in_bytes < 256 << 20:            # inputs smaller than L2 (126 MB): flush between steps
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    handle = v.Handle(x_len, h_len, "", corr)
    if args.algorithm == 0:
        os.environ["VELES_B200_NO_BATCH_CROSSOVER"] = "1"
    if args.algorithm >= 0 and args.algorithm != handle.algorithm:
        handle.close()
        handle = ForcedHandle(v, x_len, h_len, corr, args.algorithm)
    alg = handle.algorithm
    n_fft = handle.transform_length if alg != 0 else 0
    if alg == 1:
        n_fft = v.transform_length(x_len, h_len)

    def step():
        if flush is not None:
            flush.zero_()
        handle(x, h, out=y)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # clocks are sampled from before the warm-up to the end of the timed region (the timed
    # region alone can be shorter than one nvidia-smi sampling period)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    t_warm = time.perf_counter()
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    while time.perf_counter() - t_warm < args.min_warm_s:      # keep the GPU busy until samples exist
        step()
        torch.cuda.synchronize()
    barrier()
    launches0 = v.lib.simd_b200_kernel_launches()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    barrier()
    for e0, e1 in ev:
        if flush is not None:
            flush.zero_()
        e0.record()
        handle(x, h, out=y)
        e1.record()
    barrier()
    launches = v.lib.simd_b200_kernel_launches() - launches0
    # Per-kernel durations for the roofline: the same steps once more with every launch
    # bracketed by its own event pair on the same stream.  Kept out of the loop above because
    # an event record between the two launches of a step defeats their programmatic
    # dependent launch (the fused kernel starts while the spectrum kernel still runs).
    prof = {"direct": (0.0, 0), "fused": (0.0, 0), "spectrum": (0.0, 0)}
    if not args.no_profile:
        v.profile(True)
        v.profile_collect()
        for _ in range(args.steps):
            if flush is not None:
                flush.zero_()
            handle(x, h, out=y)
        torch.cuda.synchronize()
        prof = v.profile_collect()
        v.profile(False)
    clocks = sampler.stop() if rank == 0 else None
    step_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = sum(step_ms)
    t = torch.tensor([total_ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = t.item()
    ms_per_step = total_ms / args.steps
    outputs_per_step = batch * out_len * world
    value = outputs_per_step / (ms_per_step * 1e-3) / 1e9

    # ---- roofline of the dominant kernel (per launch, from its own events) ----------------
    # the batched call may cross over from the handle's single-signal choice (selector.c)
    if args.no_profile:      # no per-kernel events: charge the whole step to the dominant kernel
        kname = "direct" if v.select_algorithm(x_len, h_len, batch) == 0 else "fused"
        prof[kname] = (sum(step_ms), args.steps)
    kind = "fused" if prof["fused"][1] > 0 else "direct"
    if kind == "fused" and alg == 0:
        alg = 2
        n_fft = v.transform_length(x_len, h_len)
    k_ms, k_cnt = prof[kind]
    k_avg_ms = k_ms / max(k_cnt, 1)
    launches_per_step = max(k_cnt // args.steps, 1)
    peaks, peak_src = load_peaks()
    abytes = algorithmic_bytes(batch, x_len, h_len) / launches_per_step
    aflops = algorithmic_flops(alg, batch, x_len, h_len, n_fft) / launches_per_step
    ach_gbs = abytes / (k_avg_ms * 1e-3) / 1e9
    ach_tf = aflops / (k_avg_ms * 1e-3) / 1e12
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic_r01.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(args.workload, {}).get("dram_bytes_per_launch")
    roofline = {"bound": "hbm", "achieved": ach_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": ach_gbs / peaks["hbm_gbs"], "traffic": traffic, "peak_source": peak_src,
                "kernel": {"direct": "direct_kernel", "fused": "os_fused_kernel"}[kind],
                "kernel_ms_per_launch": k_avg_ms, "kernel_launches_per_step": launches_per_step,
                "kernel_share_of_step": min(1.0, k_avg_ms * launches_per_step / (sum(step_ms) / args.steps)),
                "timing": "per-launch event pairs on the launching stream, separate pass of the same steps",
                "algorithmic_bytes_per_launch": abytes,
                "fp32": {"achieved": ach_tf, "peak": FP32_PEAK_TFLOPS, "unit": "TFLOP/s",
                         "frac": ach_tf / FP32_PEAK_TFLOPS,
                         "peak_source": "measured FFMA microbenchmark (tools/ffma_peak.cu)",
                         "algorithmic_flops_per_launch": aflops}}

    # ---- end to end: same C-ABI call, host (pinned) buffers --------------------------------
    e2e = None
    if not args.no_e2e:
        eb = batch
        xh = torch.empty((eb, x_len), dtype=torch.float32).pin_memory()
        hh = torch.empty((h_len,), dtype=torch.float32).pin_memory()
        yh = torch.empty((eb, out_len), dtype=torch.float32).pin_memory()
        xh.copy_(x[:eb].cpu() if eb * x_len < 2 ** 28 else torch.rand((eb, x_len)) * 2 - 1)
        hh.copy_(h.cpu())
        handle(xh, hh, out=yh)            # warm-up (also sizes the staging buffers)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            handle(xh, hh, out=yh)        # returns with the result in host memory
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = tt.item()
        e2e = {"value": eb * out_len * world * args.e2e_steps / dt / 1e9, "unit": "GSamples/s",
               "h2d_bytes_per_step": 4 * (eb * x_len + h_len), "d2h_bytes_per_step": 4 * eb * out_len,
               "steps": args.e2e_steps, "ms_per_step": 1e3 * dt / args.e2e_steps,
               "result_checksum": float(yh[0, :1024].double().sum())}
        del xh, yh
    handle.close()

    base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        base, _, _ = cpu_baseline(args.workload)

    if rank == 0:
        line = {"metric": "conv1d_output_gsamples_per_s", "value": value, "unit": "GSamples/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{args.workload}: {desc}", "op": op,
                           "signals_per_gpu": batch, "x_length": x_len, "h_length": h_len,
                           "algorithm": {0: "direct", 1: "fft", 2: "overlap_save"}[alg],
                           "transform_length": n_fft,
                           "l2": ("inputs larger than L2" if flush is None
                                  else "256 MiB flush buffer written between steps"),
                           "sharding": "independent signals per rank, no collective"},
                "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
                "roofline": roofline, "cpu_baseline": base,
                "step_ms": {"min": min(step_ms), "max": max(step_ms)}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_sharded_long_signal(args, v, world, rank, local):
    """C5 on N GPUs: the 2^30-sample signal is pre-sharded (contiguous, device resident);
    each step = the (h-1)-sample halo exchange (NCCL send/recv) + the shard kernels.
    Total work is fixed, so this line reports strong scaling."""
    import torch
    import torch.distributed as dist
    from veles.simd_b200 import sharded
    op, batch, x_len, h_len, desc = WORKLOADS["c5"]
    s0, e0 = sharded.signal_range(x_len, rank, world)
    ext, x_local = sharded.alloc_shard(e0 - s0, h_len - 1, device="cuda")
    g = torch.Generator(device="cuda").manual_seed(0xC0FFEE + rank)
    x_local.copy_(torch.rand((e0 - s0,), generator=g, device="cuda") * 2 - 1)
    gh = torch.Generator(device="cuda").manual_seed(0xC0FFEE)
    h = torch.rand((h_len,), generator=gh, device="cuda") * 2 - 1
    n_out = (e0 - s0) + (h_len - 1 if rank == world - 1 else 0)
    y = torch.empty((n_out,), device="cuda")

    def step():
        sharded.filter_long_signal(x_local, h, x_len, rank, world, sharded.gpu_compute, ext=ext, out=y,
                                   overlap=bool(os.environ.get("VELES_B200_OVERLAP_HALO")))

    for _ in range(args.warmup):
        step()
    dist.barrier(); torch.cuda.synchronize()
    launches0 = v.lib.simd_b200_kernel_launches()
    e0_, e1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0_.record()
    for _ in range(args.steps):
        step()
    e1_.record()
    dist.barrier(); torch.cuda.synchronize()
    t = torch.tensor([e0_.elapsed_time(e1_)], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item() / args.steps
    if rank == 0:
        line = {"metric": "conv1d_output_gsamples_per_s", "value": (x_len + h_len - 1) / (ms * 1e-3) / 1e9,
                "unit": "GSamples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"c5: {desc}", "sharding": "contiguous shards + (h-1) halo via NCCL send/recv",
                           "l2": "inputs larger than L2"},
                "gpu_launches": int(v.lib.simd_b200_kernel_launches() - launches0),
                "e2e": None, "roofline": None, "cpu_baseline": None, "clocks": None}
        print(json.dumps(line), flush=True)
    dist.destroy_process_group()


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
