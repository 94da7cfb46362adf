#!/usr/bin/env python
"""Small deterministic target for ncu captures: `python tools/ncu_target.py gram N D [split]` runs the tcgen05 Gram 3 times,
`python tools/ncu_target.py project N D K` the tcgen05 projection 3 times, `... gram64 N D` the fp64 DMMA Gram twice,
`... project64 N D K` the fp64 DMMA projection 3 times."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

what, n, d = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
ctx = pkg.DeviceContext(0)
g = torch.Generator(device="cuda").manual_seed(1)
dt = torch.float64 if what in ("gram64", "project64") else torch.float32
X = torch.empty(n, d, device="cuda", dtype=dt)
slab = max(1, (256 << 20) // (X.element_size() * d))
for r0 in range(0, n, slab):
    X[r0:r0 + slab].normal_(generator=g)
X *= torch.linspace(1, 3, d, device="cuda", dtype=dt)
X += 0.5
shift = X[:4096].mean(0).contiguous()
if what == "fit":   # `fit N D K`: the whole PCA fit 3 times (narrow inputs: the row-folded pass)
    k = int(sys.argv[4])
    for _ in range(3):
        r = ctx.fit(X, k, center=True, algorithm="cusolver")
    ctx.sync()
    print(what, n, d, k, ctx.last_gram_path, "fold", ctx.last_fold, float(r["singular_values"][0]))
elif what in ("gram", "gram64"):
    if what == "gram":
        ctx.set_option("gram_path", 2)
        if len(sys.argv) > 4:
            ctx.set_option("gram_split", int(sys.argv[4]))
    for _ in range(3 if what == "gram" else 2):
        G = ctx.gram(X, shift)
    ctx.sync()
    print(what, n, d, ctx.last_gram_path, float(G[0, 0]))
else:
    k = int(sys.argv[4])
    Q, _ = torch.linalg.qr(torch.randn(d, k, device="cuda", generator=g, dtype=dt))
    Q = Q.T.contiguous()
    for _ in range(3):
        Z = ctx.transform(X, Q, mean=shift)
    ctx.sync()
    print(what, n, d, k, ctx.last_proj_path, float(Z[0, 0]))
