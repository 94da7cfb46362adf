"""Small end-to-end pass over every kernel family (for compute-sanitizer): pair / single-CTA / generic Gram, cluster
and grid tridiagonalisation, power path, projection kernels, streaming fit, legacy out-of-core path."""
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

ctx = pkg.DeviceContext(0)
for n, d, k, alg, center in ((3000, 256, 16, "cusolver", True), (2000, 100, 8, "cusolver", True), (1500, 320, 12, "power", False),
                             (1200, 130, 7, "cusolver", False), (900, 640, 10, "cusolver", True)):
    X = torch.randn(n, d, device="cuda") * torch.linspace(1, 3, d, device="cuda") + 0.5
    res = ctx.fit(X, k, center=center, algorithm=alg)
    Z = ctx.transform(X, res["components"], mean=res["mean"], singular_values=res["singular_values"], want_u=True)
    Xr = ctx.inverse_transform(Z[0] if isinstance(Z, tuple) else Z, res["components"], mean=res["mean"])
    ctx.sync()
    print(n, d, k, alg, float(res["singular_values"][0]), flush=True)
Xd = torch.randn(800, 96, device="cuda", dtype=torch.float64)
r = ctx.fit(Xd, 6, center=True)
ctx.sync()
ctx.partial_fit_reset()
X = torch.randn(2500, 256, device="cuda")
for a in range(0, 2500, 700):
    ctx.partial_fit(X[a:a + 700].contiguous(), center=True)
r = ctx.partial_fit_finalize(256, 8, True, torch.float32)
ctx.sync()
os.environ["DR4G_STREAM_ROWS"] = "600"
est = pkg.PCA(n_components=5, backend="gpu")
Zh = est.fit_transform(np.random.default_rng(0).standard_normal((2000, 64)).astype(np.float32))
print("ok", Zh.shape, flush=True)
