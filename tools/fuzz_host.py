"""Random-shape sweep of the drop-in call (numpy in, numpy out: pca_float / truncated_svd_float / *_double) around the staging
thresholds (inputs 12 MB, outputs 4 MB, 8 MB sub-chunks) against numpy fp64."""
import sys
from pathlib import Path
import numpy as np
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg
seed = int(sys.argv[1]) if len(sys.argv) > 1 else 0
rng = np.random.default_rng(seed)
bad = 0
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
for it in range(N):
    f64 = bool(rng.integers(0, 4) == 0)
    es = 8 if f64 else 4
    d = int(rng.integers(8, 400))
    tgt = [12 << 20, 8 << 20, 24 << 20, 4 << 20, 40 << 20][int(rng.integers(0, 5))]
    n = max(d + 1, int(tgt / (d * es) + rng.integers(-50, 51)))
    k = int(rng.integers(1, min(d, 64) + 1))
    if rng.integers(0, 2):   # outputs around 4 MB / 8 MB
        k = max(1, min(d, int(((4 << 20) if rng.integers(0, 2) else (8 << 20)) / (n * es)) + int(rng.integers(-1, 2))))
    pca = bool(rng.integers(0, 2))
    X = (rng.standard_normal((n, d)) * np.linspace(1, 3, d) + (np.linspace(-3, 3, d) if pca else 0)).astype(np.float64 if f64 else np.float32)
    cls = pkg.PCA if pca else pkg.TruncatedSVD
    est = cls(n_components=k, algorithm="cusolver", backend="gpu", dtype="float64" if f64 else None)
    Z = est.fit_transform(X)
    Xd = X.astype(np.float64); Xc = Xd - Xd.mean(0) if pca else Xd
    ev = np.sqrt(np.maximum(np.linalg.eigvalsh(Xc.T @ Xc)[::-1][:k], 0))
    e_s = float(np.max(np.abs(est.singular_values_ - ev) / ev))
    C = np.asarray(est.components_, np.float64)
    e_z = float(np.max(np.abs(Z - Xc @ C.T)) / np.max(np.abs(Xc @ C.T)))
    e_u = float(np.max(np.abs(est._U * est.singular_values_ - Z)) / np.max(np.abs(Z)))
    ok = e_s < (1e-10 if f64 else 1e-4) and e_z < (1e-11 if f64 else 2e-5) and e_u < (1e-12 if f64 else 1e-5) and Z.shape == (n, k)
    bad += not ok
    print("ok " if ok else "BAD", "f64" if f64 else "f32", "pca" if pca else "tsvd", n, d, k, f"X {n * d * es / 2**20:.1f} MB Z {n * k * es / 2**20:.1f} MB sigma {e_s:.1e} Z {e_z:.1e} U {e_u:.1e}", flush=True)
print("HOSTFUZZ", "FAILED" if bad else "PASSED", bad)
