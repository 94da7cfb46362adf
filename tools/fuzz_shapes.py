#!/usr/bin/env python
"""Random-shape sweep of fit + transform (fp32) around the dispatch boundaries (tiny inputs, row folding, padded chunk copies,
pair / single-CTA Gram, k slices) against fp64 references formed on the GPU: sigma at 1e-4, scores at 2e-5 of max|Z|."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 0
cases = int(sys.argv[2]) if len(sys.argv) > 2 else 60
f64 = len(sys.argv) > 3 and sys.argv[3] == "f64"
dt = torch.float64 if f64 else torch.float32
TS, TZ = (1e-10, 1e-11) if f64 else (1e-4, 2e-5)
rng = np.random.default_rng(seed)
ctx = pkg.DeviceContext(0)
bad = 0
for it in range(cases):
    kind = rng.integers(0, 5)
    if kind == 0:   # around the tiny threshold
        d = int(rng.integers(1, 300)); n = max(2, int(262144 // d + rng.integers(-3, 4)))
    elif kind == 1:  # around the fold threshold
        d = int(rng.integers(1, 129)); n = max(2, int((1 << 24) // d + rng.integers(-40, 41)))
    elif kind == 2:  # folded, ragged tails
        d = int(rng.integers(1, 129)); n = int((1 << 24) // d * rng.uniform(1.0, 2.5))
    elif kind == 3:  # wide, odd d
        d = int(rng.integers(129, 1400)); n = int(rng.integers(300, 60000))
    else:            # anything
        d = int(rng.integers(1, 700)); n = int(rng.integers(1, 200000))
    k = int(rng.integers(1, min(d, n, 150) + 1))
    center = bool(rng.integers(0, 2))
    g = torch.Generator(device="cuda").manual_seed(int(rng.integers(1 << 30)))
    X = (torch.randn(n, d, device="cuda", generator=g, dtype=dt) * torch.linspace(1, 3, d, device="cuda", dtype=dt) + (torch.linspace(-4, 4, d, device="cuda", dtype=dt) if center else 0.0)).contiguous()
    if f64 and n * d > 3e7:
        continue
    try:
        r = ctx.fit(X, k, center=center, algorithm="cusolver")
        ctx.sync()
        fold, path = ctx.last_fold, ctx.last_gram_path
        Xd = X.double()
        mu = Xd.mean(0) if center else torch.zeros(d, device="cuda", dtype=torch.float64)
        G = (Xd - mu).T @ (Xd - mu)
        ev = torch.linalg.eigvalsh(G).flip(0)[:k].clamp(min=0).sqrt()
        s = r["singular_values"].double()
        big = ev > 1e-3 * ev[0]
        es = float(((s - ev).abs() / ev.clamp(min=1e-300))[big].max()) if bool(big.any()) else 0.0
        em = float((r["mean"].double() - mu).abs().max()) if center else 0.0
        Z = ctx.transform(X, r["components"], mean=r["mean"] if center else None)
        ctx.sync()
        Zr = (Xd - mu) @ r["components"].double().T
        ez = float((Z.double() - Zr).abs().max() / Zr.abs().max().clamp(min=1e-30))
        C = r["components"].double()
        eo = float((C @ C.T - torch.eye(k, device="cuda", dtype=torch.float64)).abs().max())
        ok = es < TS and ez < TZ and em < TS and eo < (1e-9 if f64 else 5e-3) and bool(torch.isfinite(s).all())
        tag = "ok " if ok else "BAD"
    except Exception as e:  # noqa: BLE001
        ok, tag, es, ez, em, eo, fold, path = False, "EXC", 0, 0, 0, 0, -1, str(e)[:120]
    bad += not ok
    print(f"{tag} n={n} d={d} k={k} center={center} gram={path} fold={fold} proj={ctx.last_proj_path} sigma={es:.1e} Z={ez:.1e} mean={em:.1e} orth={eo:.1e}", flush=True)
    del X
print("FUZZ", "FAILED" if bad else "PASSED", bad, "of", cases)
