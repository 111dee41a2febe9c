import os, sys, statistics, torch
sys.path.insert(0, "/root/repo")
import veles.simd_b200 as v
v.use_torch_stream()
B, xl = 256, 262144
x = torch.rand((B, xl), device="cuda") * 2 - 1
for hl in (3000, 4096, 4500, 5000, 5500, 6000, 7000, 8192):
    h = torch.rand((hl,), device="cuda") * 2 - 1
    y = torch.empty((B, xl + hl - 1), device="cuda")
    res = {}
    for mode in ("0", "1"):
        os.environ["VELES_B200_REAL"] = mode
        os.environ["VELES_B200_FORCE_N"] = "16384"
        with v.Handle(xl, hl, "_overlap_save") as inner:
            hd = v.ConvolutionHandle()
            hd.algorithm, hd.x_length, hd.h_length = 2, xl, hl
            hd.handle.os = inner.handle
            fn = lambda: v.lib.convolve_batch(hd, x.data_ptr(), xl, h.data_ptr(), y.data_ptr(), xl + hl - 1, B)
            for _ in range(3): fn()
            torch.cuda.synchronize()
            ts = []
            for _ in range(7):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
            res[mode] = statistics.median(ts)
    print(f"h={hl}: paired {res['0']:.0f} us, real-packed {res['1']:.0f} us, ratio {res['1']/res['0']:.3f}")
