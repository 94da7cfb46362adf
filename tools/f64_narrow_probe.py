"""fp64 fits of narrow inputs: stage times and the rate of the streamed pass."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg
ctx = pkg.DeviceContext(0)
for n, d, k in [(1 << 22, 64, 8), (1 << 23, 32, 8), (1 << 21, 128, 16), (1 << 24, 16, 4), (1 << 21, 130, 16), (1 << 21, 256, 16)]:
    g = torch.Generator(device="cuda").manual_seed(n + d)
    X = torch.randn(n, d, device="cuda", generator=g, dtype=torch.float64) * torch.linspace(1, 3, d, device="cuda", dtype=torch.float64) + 0.5
    for _ in range(3):
        r = ctx.fit(X, k, center=True, algorithm="cusolver"); ctx.sync()
    st = ctx.stage_times()
    Z = ctx.transform(X, r["components"], mean=r["mean"]); ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ctx.transform(X, r["components"], mean=r["mean"], out=Z); e1.record(); torch.cuda.synchronize()
    print(f"n={n} d={d} fp64: gram {st['gram']:.3f} ms = {n * d * 8 / st['gram'] / 1e6:.0f} GB/s ({ctx.last_gram_path}); transform {e0.elapsed_time(e1):.3f} ms = {n * (d + k) * 8 / e0.elapsed_time(e1) / 1e6:.0f} GB/s ({ctx.last_proj_path})", flush=True)
    del X, Z
