#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nproc
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_streaming.py tests/test_gpu_baseline_configs.py -x -q -m gpu 2>&1 | tail -3
cat > /tmp/pg.py <<'PY'
import sys, time, os
import numpy as np
sys.path.insert(0, '.')
import dimreduce4gpu_b200 as pkg
rng = np.random.default_rng(0)
for n, d, k in [(524288, 512, 32), (1000000, 256, 32)]:
    X = (rng.standard_normal((n, d), dtype=np.float32) * np.linspace(1, 3, d, dtype=np.float32) + np.linspace(-5, 5, d, dtype=np.float32))
    est = pkg.PCA(n_components=k, backend='gpu')
    ts = []
    for i in range(4):
        t0 = time.perf_counter(); Z = est.fit_transform(X); ts.append(time.perf_counter() - t0)
    print(os.environ.get('DR4G_NO_STAGING', 'staged'), os.environ.get('DR4G_STAGE_THREADS', '-'), n, d, k, 'ms', [round(t * 1e3, 1) for t in ts], 'rows/s', f"{n / min(ts[1:]):.3e}", float(est.singular_values_[0]))
PY
DR4G_NO_STAGING=1 python /tmp/pg.py
python /tmp/pg.py
for t in 2 4 8 12; do DR4G_STAGE_THREADS=$t python /tmp/pg.py | head -1; done
