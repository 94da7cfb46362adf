"""Probe: where the reconstruction error of the rank-deficient 200 x 80 reference case comes from."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests" / "golden"))
import inputs as I  # noqa: E402
import dimreduce4gpu_b200 as pkg  # noqa: E402

X = I.comprehensive_matrix(200, 80, 16, True, np.float32, 1)
k = 16
X64 = X.astype(np.float64); mu = X64.mean(0); Xc = X64 - mu
rel = lambda A, B: np.linalg.norm(A - B) / np.linalg.norm(B)
w, V = np.linalg.eigh(Xc.T @ Xc); C = V[:, ::-1][:, :k].T
for alg in ("power", "cusolver"):
    m = pkg.PCA(n_components=k, backend="gpu", algorithm=alg, n_iter=5, random_state=1)
    Z = m.fit_transform(X)
    Cg = np.asarray(m.components_, np.float64)
    print(alg, "recon err", rel(Z.astype(np.float64) @ Cg, Xc))
    print("  mean err", np.abs(m.mean_ - mu).max(), "sigma rel", np.abs(m.singular_values_ - np.sqrt(w[::-1][:k])).max() / np.sqrt(w[-1]))
    P = C.T @ C
    print("  components outside true span", np.linalg.norm(Cg - Cg @ P), "orth", np.abs(Cg @ Cg.T - np.eye(k)).max())
    print("  Z vs fp64 (X-mu)Cg^T", rel(Z.astype(np.float64), Xc @ Cg.T), " recon with fp64 Z", rel((Xc @ Cg.T) @ Cg, Xc))
    ctx = pkg.truncated_svd._device_context(0)
    print("  gram path", ctx.last_gram_path, "proj path", ctx.last_proj_path)
