#!/usr/bin/env python
"""Probe: small-d Gram as it is dispatched today vs the same rows viewed as (n/f) x (f d) ("row folding") on the pair kernel."""
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


for n, d in [(1 << 22, 128), (1 << 23, 64), (1 << 24, 32), (1 << 25, 16), (1 << 26, 8), (1 << 22, 100), (1 << 22, 96), (1 << 23, 48), (1 << 20, 128), (100000, 64)]:
    g = torch.Generator(device="cuda").manual_seed(n + d)
    X = torch.randn(n, d, device="cuda", generator=g) * torch.linspace(1, 3, d, device="cuda") + 0.5
    shift = X[:4096].mean(0).contiguous()
    G0 = ctx.gram(X, shift).clone(); ctx.sync()
    p0 = ctx.last_gram_path
    t0 = timed(lambda: ctx.gram(X, shift))
    f = 256 // d
    while (f * d) % 4:
        f -= 1
    n2 = n // f
    X2 = X[: n2 * f].view(n2, f * d)
    s2 = shift.repeat(f).contiguous()
    G2 = ctx.gram(X2, s2).clone(); ctx.sync()
    p2 = ctx.last_gram_path
    t2 = timed(lambda: ctx.gram(X2, s2))
    Gf = sum(G2[i * d:(i + 1) * d, i * d:(i + 1) * d] for i in range(f))
    if n2 * f < n:
        Y = (X[n2 * f:] - shift).double()
        Gf = Gf + Y.T @ Y
    dg = torch.sqrt(G0.diagonal())
    err = float(((Gf - G0).abs() / (dg[:, None] * dg[None, :])).max())
    print(f"n={n} d={d}: now path={p0} {t0:.3f} ms = {n * d * 4 / t0 / 1e6:.0f} GB/s | folded f={f} d'={f * d} path={p2} {t2:.3f} ms = {n * d * 4 / t2 / 1e6:.0f} GB/s | diff {err:.1e}", flush=True)
    del X, X2, G0, G2
