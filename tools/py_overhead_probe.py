import sys, time
from pathlib import Path
import numpy as np
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg
from dimreduce4gpu_b200._backend import select_backend
from dimreduce4gpu_b200.lib_dimreduce4gpu import _load_pca_lib
from dimreduce4gpu_b200._native import last_error
import cProfile, pstats
def t(f, reps=200):
    f(); t0 = time.perf_counter()
    for _ in range(reps): f()
    return (time.perf_counter() - t0) / reps * 1e6
X = np.random.default_rng(0).standard_normal((10000, 128)).astype(np.float32)
print("select_backend us", t(lambda: select_backend("gpu")))
print("_load_pca_lib us", t(lambda: _load_pca_lib("float")))
print("last_error us", t(last_error))
print("np.zeros(7) us", t(lambda: [np.zeros(s, np.float32) for s in ((32, 128), (32,), (10000, 32), (10000, 32), (32,), (32,), (128,))]))
print("ascontiguous us", t(lambda: np.ascontiguousarray(X, dtype=np.float32)))
est = pkg.PCA(n_components=32, backend="gpu")
est.fit_transform(X)
pr = cProfile.Profile(); pr.enable()
for _ in range(50): est.fit_transform(X)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(12)
