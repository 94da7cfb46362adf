#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
timeout 600 python -m pytest tests/test_gpu_project_tc.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -4
timeout 300 python - <<'PY'
import sys, torch
sys.path.insert(0, '.')
import dimreduce4gpu_b200 as pkg
for path in (0, 1):
    ctx = pkg.DeviceContext(0); ctx.set_option('proj_path', path)
    for n, d, k in [(1000000,256,2),(1000000,256,10),(1000000,784,50),(1000000,256,32)]:
        g = torch.Generator(device='cuda').manual_seed(n+d)
        X = torch.randn(n, d, device='cuda', generator=g)
        Q, _ = torch.linalg.qr(torch.randn(d, k, device='cuda', generator=g)); Q = Q.T.contiguous()
        mean = X[:4096].mean(0).contiguous()
        Z = ctx.transform(X, Q, mean=mean); torch.cuda.synchronize()
        ts = []
        for _ in range(7):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ctx.transform(X, Q, mean=mean, out=Z); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[3]
        print(f"proj path={ctx.last_proj_path} n={n} d={d} k={k} ms={ms:.3f} GB/s={n*(d+k)*4/ms/1e6:.0f}", flush=True)
        del X, Z
PY
