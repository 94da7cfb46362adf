"""GPU probe: fp64 fit (generic DMMA Gram + eigensolver) stage times for one shape (n, d, k)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

n, d, k = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (200000, 4096, 100)))
ctx = pkg.DeviceContext(0)
g = torch.Generator(device="cuda").manual_seed(1)
X = torch.randn(n, d, device="cuda", dtype=torch.float64, generator=g) * torch.linspace(1, 3, d, device="cuda", dtype=torch.float64) + 0.5
out = None
for i in range(2):
    out = ctx.fit(X, k, center=True, algorithm="cusolver", out=out)
    torch.cuda.synchronize()
    print({kk: round(v, 3) for kk, v in ctx.stage_times().items()}, ctx.last_gram_path, flush=True)
Xc = X - X.mean(0)
s = torch.linalg.svdvals(Xc[:, :]) if n * d <= 2e8 else None
if s is not None:
    rel = ((out["singular_values"] - s[:k]).abs() / s[:k]).max().item()
    print("sigma rel err vs torch.linalg.svdvals:", rel)
