"""Wall-clock of the drop-in call on small inputs (what the reference's bench/run_benchmarks.py times: fit_transform of a numpy array)."""
import sys, time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402

rng = np.random.default_rng(0)
for name, cls, n, d, k, alg in [("4x3", pkg.PCA, 4, 3, 2, "cusolver"), ("c1", pkg.TruncatedSVD, 10000, 100, 10, "power"),
                                ("r1 pca", pkg.PCA, 10000, 128, 32, "cusolver"), ("r1 tsvd", pkg.TruncatedSVD, 10000, 128, 32, "power"),
                                ("r2 pca", pkg.PCA, 50000, 256, 64, "cusolver"), ("r2 tsvd", pkg.TruncatedSVD, 50000, 256, 64, "power"),
                                ("c2/10", pkg.PCA, 100000, 256, 32, "cusolver")]:
    X = rng.standard_normal((n, d)).astype(np.float32)
    est = cls(n_components=k, algorithm=alg, backend="gpu", random_state=0)
    est.fit_transform(X)
    ts = []
    for _ in range(15):
        t0 = time.perf_counter(); Z = est.fit_transform(X); ts.append(time.perf_counter() - t0)
    ts.sort()
    t1 = []
    for _ in range(15):
        e2 = cls(n_components=k, algorithm=alg, backend="gpu", random_state=0)
        t0 = time.perf_counter(); e2.fit_transform(X); t1.append(time.perf_counter() - t0)
    t1.sort()
    print(f"{name:8s} {n}x{d} k={k}: fit_transform median {ts[7] * 1e3:.3f} ms (min {ts[0] * 1e3:.3f}); new estimator each call {t1[7] * 1e3:.3f} ms; PCIe floor {(n * d + 2 * n * k) * 4 / 55e9 * 1e3:.3f} ms", flush=True)
