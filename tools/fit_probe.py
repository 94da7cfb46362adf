"""GPU probe: stage times of DeviceContext.fit for one shape (n,d,k), optionally with the Gram pipeline trace."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

n, d, k = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (1000000, 256, 32)))
center = (sys.argv[4] != "0") if len(sys.argv) > 4 else True
ctx = pkg.DeviceContext(0)
X = torch.randn(n, d, device="cuda") * torch.linspace(1, 3, d, device="cuda") + (torch.linspace(-5, 5, d, device="cuda") if center else 0.0)
out = None
for i in range(4):
    out = ctx.fit(X, k, center=center, algorithm="cusolver", out=out)
    torch.cuda.synchronize()
    print({kk: round(v, 4) for kk, v in ctx.stage_times().items()}, ctx.last_gram_path, flush=True)
