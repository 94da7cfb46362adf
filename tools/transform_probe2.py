"""K7 timing without the allocation of Z inside the timed region (out= reused), several shapes."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

ctx = pkg.DeviceContext(0)
for n, d, k in ((1000000, 256, 32), (4194304, 512, 32), (4194304, 1024, 64), (1000000, 784, 50), (8000000, 128, 16), (1000000, 256, 128)):
    g = torch.Generator(device="cuda").manual_seed(n + d)
    X = torch.randn(n, d, device="cuda", generator=g) + 1.0
    Q, _ = torch.linalg.qr(torch.randn(d, k, device="cuda", generator=g, dtype=torch.float64))
    Q = Q.T.contiguous().float()
    mean = X[:100000].mean(0).contiguous()
    Z = ctx.transform(X, Q, mean=mean)
    ctx.sync()
    ts = []
    for _ in range(9):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctx.transform(X, Q, mean=mean, out=Z); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[4]
    print(f"n={n} d={d} k={k}: transform {ms:.3f} ms, {(d + k) * 4 * n / ms / 1e6:.0f} GB/s algorithmic = {(d + k) * 4 * n / ms / 1e6 / 6548:.2f} of HBM copy", flush=True)
    del X, Z
