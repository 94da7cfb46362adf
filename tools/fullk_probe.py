"""k = d fits (all components) at a few sizes, fp32 and fp64, against fp64 eigenvalues formed on the GPU."""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg
ctx = pkg.DeviceContext(0)
bad = 0
for dt, tol in ((torch.float32, 1e-4), (torch.float64, 1e-10)):
    for n, d in [(900, 300), (2500, 700), (4000, 1100), (200, 500), (9000, 2100)]:
        k = min(n, d)
        g = torch.Generator(device="cuda").manual_seed(n + d)
        X = (torch.randn(n, d, device="cuda", generator=g, dtype=dt) * torch.linspace(1, 3, d, device="cuda", dtype=dt) + 0.5).contiguous()
        r = ctx.fit(X, k, center=True, algorithm="cusolver"); ctx.sync()
        Xd = X.double(); Xc = Xd - Xd.mean(0)
        ev = torch.linalg.eigvalsh(Xc.T @ Xc).flip(0)[:k].clamp(min=0).sqrt()
        s = r["singular_values"].double()
        big = ev > 1e-3 * ev[0]
        es = float(((s - ev).abs() / ev)[big].max())
        C = r["components"].double()[: int(big.sum())]
        eo = float((C @ C.T - torch.eye(C.shape[0], device="cuda", dtype=torch.float64)).abs().max())
        Z = ctx.transform(X, r["components"], mean=r["mean"]); ctx.sync()
        Zr = Xc @ r["components"].double().T
        ez = float((Z.double() - Zr).abs().max() / Zr.abs().max())
        ok = es < tol and eo < (1e-9 if dt == torch.float64 else 5e-3) and ez < (1e-11 if dt == torch.float64 else 2e-5)
        bad += not ok
        print("ok " if ok else "BAD", dt, n, d, k, f"sigma {es:.1e} orth {eo:.1e} Z {ez:.1e}", ctx.stage_times()["eigensolve"], flush=True)
print("FULLK", "FAILED" if bad else "PASSED")
