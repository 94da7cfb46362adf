import sys, time
from pathlib import Path
import numpy as np
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg
rng = np.random.default_rng(0)
for n, d, k in [(50000, 256, 64), (100000, 256, 32), (1000000, 256, 32)]:
    X = rng.standard_normal((n, d), dtype=np.float32)
    est = pkg.PCA(n_components=k, backend="gpu", random_state=0)
    for _ in range(4):
        t0 = time.perf_counter(); est.fit_transform(X); print(f"-- {n}x{d}: {1e3 * (time.perf_counter() - t0):.3f} ms wall", file=sys.stderr)
