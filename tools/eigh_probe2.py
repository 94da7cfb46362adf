"""GPU probe: top-k eigensolver (tridiagonal route) vs torch.linalg.eigh in fp64, with timing."""
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

ctx = pkg.DeviceContext(0)
shapes = [(1, 1), (2, 2), (3, 2), (5, 5), (33, 7), (100, 10), (256, 32), (256, 256), (512, 32), (1024, 64)]
if len(sys.argv) > 1:
    shapes = [tuple(map(int, s.split(","))) for s in sys.argv[1].split(";")]
for kind in ("aniso", "flat", "lowrank"):
    for d, k in shapes:
        g = torch.Generator(device="cuda").manual_seed(d * 31 + k)
        n = 4 * d + 50
        X = torch.randn(n, d, device="cuda", dtype=torch.float64, generator=g)
        if kind == "aniso":
            X = X * torch.linspace(1, 3, d, device="cuda", dtype=torch.float64)
        elif kind == "lowrank":
            r = max(1, d // 3)
            X = X[:, :r] @ torch.randn(r, d, device="cuda", dtype=torch.float64, generator=g)
        A = X.T @ X
        w, Vt = ctx.eigh_topk(A, k)
        ctx.sync()
        wr, Vr = torch.linalg.eigh(A)
        wr = wr.flip(0)[:k]
        scale = wr[0].abs().item() + 1e-300
        werr = ((w - wr).abs().max() / scale).item()
        res = ((A @ Vt.T - Vt.T * w).norm(dim=0).max() / scale).item()
        orth = (Vt @ Vt.T - torch.eye(k, device="cuda", dtype=torch.float64)).abs().max().item()
        ts = []
        for _ in range(3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ctx.eigh_topk(A, k)
            ctx.sync()
            ts.append(time.perf_counter() - t0)
        print(f"{kind:8s} d={d:5d} k={k:4d} eig_err={werr:.2e} resid={res:.2e} orth={orth:.2e} ms={min(ts) * 1e3:8.3f}", flush=True)
