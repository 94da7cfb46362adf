"""Where the drop-in call's time goes on the reference's bench shape (50,000 x 256, k = 64): fresh vs reused (touched) output arrays,
staging on / off, and the pieces of the wrapper."""
import ctypes, os, sys, time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402
from dimreduce4gpu_b200.lib_dimreduce4gpu import _load_pca_lib, params  # noqa: E402

def ptr(a): return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
rng = np.random.default_rng(0)
for n, d, k in [(50000, 256, 64), (100000, 256, 32), (10000, 128, 32)]:
    X = rng.standard_normal((n, d)).astype(np.float32)
    f = _load_pca_lib("float")
    p = params(); p.X_n, p.X_m, p.k = n, d, k; p.algorithm = b"cusolver"; p.n_iter = 5; p.random_state = 0; p.tol = 1e-5; p.verbose = 0; p.gpu_id = 0; p.whiten = False
    def alloc():
        return [np.zeros(s, np.float32) for s in ((k, d), (k,), (n, k), (n, k), (k,), (k,), (d,))]
    def call(o):
        Q, w, U, Xt, ev, evr, mean = o
        f(ptr(X), ptr(Q), ptr(w), ptr(U), ptr(Xt), ptr(ev), ptr(evr), ptr(mean), p)
    o = alloc(); call(o); call(o)
    ta, tf, tw = [], [], []
    for _ in range(11):
        t0 = time.perf_counter(); o2 = alloc(); t1 = time.perf_counter(); call(o2); t2 = time.perf_counter(); call(o); t3 = time.perf_counter()
        ta.append(t1 - t0); tf.append(t2 - t1); tw.append(t3 - t2)
    med = lambda v: sorted(v)[len(v) // 2] * 1e3
    print(f"{n}x{d} k={k} staging={'off' if os.environ.get('DR4G_NO_STAGING') else 'on'}: np.zeros {med(ta):.3f} ms | native call, fresh outputs {med(tf):.3f} ms | native call, touched outputs {med(tw):.3f} ms", flush=True)
