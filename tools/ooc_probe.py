"""GPU probe: pca_float through the legacy host-pointer symbol, device-resident path vs the out-of-core path
(DR4G_STREAM_ROWS), pinned host buffers, wall clock around the synchronous call."""
import ctypes
import os
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from dimreduce4gpu_b200.lib_dimreduce4gpu import _load_pca_lib, params  # noqa: E402

n, d, k = 1_000_000, 256, 32
Xh = torch.empty(n, d, dtype=torch.float32, pin_memory=True)
Xh.copy_(torch.randn(n, d, device="cuda") * torch.linspace(1, 3, d, device="cuda") + 1.0)
pin = lambda *s: torch.zeros(*s, dtype=torch.float32, pin_memory=True)  # noqa: E731
Q, w, U, Xt, ev, evr, mean = pin(k, d), pin(k), pin(n, k), pin(n, k), pin(k), pin(k), pin(d)
p = params()
p.X_n, p.X_m, p.k = n, d, k
p.algorithm = b"cusolver"
p.n_iter, p.random_state, p.tol, p.verbose, p.gpu_id, p.whiten = 5, 42, 1e-5, 0, 0, False
fp = lambda t: ctypes.cast(t.data_ptr(), ctypes.POINTER(ctypes.c_float))  # noqa: E731
fn = _load_pca_lib()
args = [fp(Xh), fp(Q), fp(w), fp(U), fp(Xt), fp(ev), fp(evr), fp(mean), p]
ref_w = None
for rows in (0, 32768, 131072, 262144, 524288):
    if rows:
        os.environ["DR4G_STREAM_ROWS"] = str(rows)
    else:
        os.environ.pop("DR4G_STREAM_ROWS", None)
    ts = []
    for i in range(4):
        t0 = time.perf_counter()
        fn(*args)
        if i:
            ts.append(time.perf_counter() - t0)
    if ref_w is None:
        ref_w = w.clone()
    err = float(((w - ref_w).abs() / ref_w).max())
    print(f"{'resident' if not rows else f'out-of-core, {rows} rows/chunk':34s}: {sorted(ts)[1] * 1e3:7.2f} ms   "
          f"(H2D {n * d * 4 * (2 if rows else 1) / 1e9:.2f} GB, D2H {2 * n * k * 4 / 1e9:.2f} GB; sigma rel diff {err:.1e})", flush=True)
