#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
cat > /tmp/pg.py <<'PY'
import sys, time, os
import numpy as np
sys.path.insert(0, '.')
import dimreduce4gpu_b200 as pkg
rng = np.random.default_rng(0)
for n, d, k in [(50000, 256, 64), (524288, 512, 32), (1000000, 256, 32)]:
    X = (rng.standard_normal((n, d), dtype=np.float32) * np.linspace(1, 3, d, dtype=np.float32) + np.linspace(-5, 5, d, dtype=np.float32))
    est = pkg.PCA(n_components=k, backend='gpu')
    ts = []
    for i in range(8):
        t0 = time.perf_counter(); Z = est.fit_transform(X); ts.append(time.perf_counter() - t0)
    ts = sorted(ts[1:])
    print(os.environ.get('DR4G_STAGE_THREADS', 'default'), n, d, k, 'median ms', round(ts[3] * 1e3, 2), 'min', round(ts[0] * 1e3, 2), flush=True)
PY
for t in default 4 8 12 16; do if [ $t = default ]; then python /tmp/pg.py; else DR4G_STAGE_THREADS=$t python /tmp/pg.py; fi; done
python tools/small_e2e_probe.py
