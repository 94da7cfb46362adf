#!/usr/bin/env python
"""Timing of the tcgen05 Gram on the BASELINE shapes: `python tools/gram_bench.py TAG [split] [shapes n,d;n,d]` (median of 7, CUDA events)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

tag = sys.argv[1]
split = int(sys.argv[2]) if len(sys.argv) > 2 else 2
shapes = sys.argv[3] if len(sys.argv) > 3 else "1000000,256;4194304,512;2097152,1024;1000000,768"
ctx = pkg.DeviceContext(0)
ctx.set_option("gram_path", 2)
ctx.set_option("gram_split", split)
import os
if os.environ.get("FLUSH_ROWS"):
    ctx.set_option("flush_rows", int(os.environ["FLUSH_ROWS"]))
for spec in shapes.split(";"):
    n, d = map(int, spec.split(","))
    g = torch.Generator(device="cuda").manual_seed(n * 7 + d)
    X = torch.randn(n, d, device="cuda", generator=g) * torch.linspace(1, 3, d, device="cuda") + 0.5
    shift = X[:4096].mean(0).contiguous()
    G = ctx.gram(X, shift)
    ctx.sync()
    m = min(n, 200000)
    Y = (X[:m] - shift).double()
    R = Y.T @ Y
    Gm = ctx.gram(X[:m], shift).clone()
    ctx.sync()
    dg = torch.sqrt(R.diagonal())
    err = float(((Gm - R).abs() / (dg[:, None] * dg[None, :])).max())
    ts = []
    for _ in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctx.gram(X, shift); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[3]
    print(f"{tag} split={split} n={n} d={d} path={ctx.last_gram_path} err({m} rows)={err:.2e} ms={ms:.3f} "
          f"alg_TF/s={2 * n * d * d / ms / 1e9:.1f} GB/s={n * d * 4 / ms / 1e6:.0f}", flush=True)
    del X, G, Y, R
