"""GPU probe: K7 projection (DeviceContext.transform) time and achieved HBM GB/s, CUDA events on torch's stream."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

ctx = pkg.DeviceContext(0)
for n, d, k in ((1000000, 256, 32), (4000000, 512, 32), (2000000, 1024, 64), (1000000, 100, 10)):
    X = torch.randn(n, d, device="cuda") + 1.0
    res = ctx.fit(X, k, center=True, algorithm="cusolver")
    Z = ctx.transform(X, res["components"], mean=res["mean"])
    ctx.sync()
    ref = (X.double() - res["mean"].double()) @ res["components"].double().T
    err = ((Z.double() - ref).abs().max() / ref.abs().max()).item()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        Z = ctx.transform(X, res["components"], mean=res["mean"])
        ctx.sync()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[2]
    print(f"n={n} d={d} k={k}: transform {ms:.3f} ms, {(d + k) * 4 * n / ms / 1e6:.0f} GB/s algorithmic, rel err {err:.2e}", flush=True)
