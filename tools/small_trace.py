import sys, time
from pathlib import Path
import numpy as np
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg
rng = np.random.default_rng(0)
for n, d, k, cls, alg in [(4, 3, 2, pkg.PCA, "cusolver"), (10000, 100, 10, pkg.TruncatedSVD, "power"), (10000, 128, 32, pkg.PCA, "cusolver")]:
    X = rng.standard_normal((n, d)).astype(np.float32)
    est = cls(n_components=k, algorithm=alg, backend="gpu", random_state=0)
    for _ in range(4):
        t0 = time.perf_counter(); est.fit_transform(X); print(f"-- {n}x{d}: {1e3 * (time.perf_counter() - t0):.3f} ms wall", file=sys.stderr)
