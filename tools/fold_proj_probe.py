#!/usr/bin/env python
"""Probe: K7 on narrow inputs as dispatched today vs the same rows folded f-fold with block-diagonal components."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

ctx = pkg.DeviceContext(0)


def timed(fn, reps=7):
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


for n, d, k in [(1 << 22, 128, 32), (1 << 23, 64, 16), (1 << 24, 32, 8), (1 << 25, 16, 4), (1 << 26, 8, 2), (1 << 22, 100, 10), (1 << 24, 32, 2), (1 << 23, 48, 6)]:
    g = torch.Generator(device="cuda").manual_seed(n + d)
    X = torch.randn(n, d, device="cuda", generator=g) * torch.linspace(1, 3, d, device="cuda") + 0.5
    Q, _ = torch.linalg.qr(torch.randn(d, k, device="cuda", generator=g, dtype=torch.float64))
    Q = Q.T.contiguous().float()
    mean = X[:100000].mean(0).contiguous()
    Z0 = ctx.transform(X, Q, mean=mean); ctx.sync()
    p0 = ctx.last_proj_path
    t0 = timed(lambda: ctx.transform(X, Q, mean=mean, out=Z0))
    f = min(256 // d, 128 // k)
    while (f * d) % 4:
        f -= 1
    n2 = n // f
    Qf = torch.zeros(f * k, f * d, device="cuda")
    for gidx in range(f):
        Qf[gidx * k:(gidx + 1) * k, gidx * d:(gidx + 1) * d] = Q
    X2 = X[: n2 * f].view(n2, f * d)
    m2 = mean.repeat(f).contiguous()
    Z2 = ctx.transform(X2, Qf, mean=m2); ctx.sync()
    p2 = ctx.last_proj_path
    t2 = timed(lambda: ctx.transform(X2, Qf, mean=m2, out=Z2))
    err = float((Z2.view(-1, k) - Z0[: n2 * f]).abs().max() / Z0.abs().max())
    B = n * (d + k) * 4
    print(f"n={n} d={d} k={k}: now path={p0} {t0:.3f} ms = {B / t0 / 1e6:.0f} GB/s | folded f={f} d'={f * d} k'={f * k} path={p2} {t2:.3f} ms = {B / t2 / 1e6:.0f} GB/s | diff {err:.1e}", flush=True)
    del X, X2, Z0, Z2
