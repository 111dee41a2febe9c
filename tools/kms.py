import json, sys
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
r = d["roofline"]
print(sys.argv[1] if len(sys.argv) > 1 else "", d["config"]["workload"][:2], f'{d["value"]:.1f} GS/s step {d["ms_per_step"]:.4f} ms kernel {r["kernel_ms_per_launch"]:.4f} ms hbm {r["frac"]:.3f} fp32 {r["fp32"]["frac"]:.3f}')
