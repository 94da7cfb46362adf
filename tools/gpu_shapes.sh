#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
timeout 600 python - <<'PY'
import sys, torch, time
sys.path.insert(0, '.')
import dimreduce4gpu_b200 as pkg
ctx = pkg.DeviceContext(0)
def ev(fn, reps=5):
    fn(); torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts)//2]
for dt in (torch.float32, torch.float64):
    for n, d, k in [(1000000, 50, 10), (1000000, 99, 10), (500000, 1001, 32), (1000000, 30, 5), (1000000, 254, 32), (200000, 1024, 200), (1000000, 128, 16), (1000000, 64, 8)]:
        if dt == torch.float64 and n * d > 3e8: n //= 2
        X = torch.randn(n, d, device='cuda', dtype=dt) * torch.linspace(1, 3, d, device='cuda', dtype=dt)
        fit = ev(lambda: ctx.fit(X, k, center=True), 3)
        st = ctx.stage_times(); gp = ctx.last_gram_path
        o = ctx.fit(X, k, center=True)
        tr = ev(lambda: ctx.transform(X, o['components'], mean=o['mean']), 3); pp = ctx.last_proj_path
        Z = ctx.transform(X, o['components'], mean=o['mean'])
        inv = ev(lambda: ctx.inverse_transform(Z, o['components'], mean=o['mean']), 3)
        es = X.element_size()
        print(f"{str(dt)[6:]:8s} n={n:8d} d={d:5d} k={k:4d} fit {fit:8.3f} ms (gram {st['gram']:.3f} {gp}, {n*d*es/st['gram']/1e6:6.0f} GB/s; eig {st['eigensolve']:.3f}) transform {tr:7.3f} ms {pp} {n*(d+k)*es/tr/1e6:6.0f} GB/s  inverse {inv:7.3f} ms {n*(d+k)*es/inv/1e6:6.0f} GB/s", flush=True)
        del X, Z
PY
