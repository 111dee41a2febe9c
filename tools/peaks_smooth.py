import sys, torch
sys.path.insert(0, '/root/repo')
import veles.simd_b200 as v
v.use_torch_stream()
n = 1 << 28
t = torch.arange(n, device='cuda', dtype=torch.float32)
x = torch.sin(t * 1e-3)
out = torch.zeros((200000, 2), dtype=torch.int32, device='cuda')
for _ in range(3): c = v.detect_peaks_device(x, out, 3)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): c = v.detect_peaks_device(x, out, 3)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
print(f"smooth series 2^28 samples, {c} extrema: {ms:.3f} ms = {n/ms/1e6:.0f} GSamples/s = {4*n/ms/1e6:.0f} GB/s of input")
