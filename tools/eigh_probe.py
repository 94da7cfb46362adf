"""GPU probe: time/accuracy of dr4g_eigh_double vs sweep cap (development aid, not a test)."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from dimreduce4gpu_b200 import DeviceContext

ctx = DeviceContext(0)
for d in [int(a) for a in sys.argv[1:]] or [100, 256, 512, 1024]:
    rng = np.random.default_rng(d)
    X = rng.standard_normal((4 * d, d)) * np.linspace(1, 3, d)
    A = torch.from_numpy(X.T @ X).cuda()
    wref = np.linalg.eigvalsh(A.cpu().numpy())
    for cap in (2, 4, 6, 8, 10, 12, 30):
        ctx.set_option("jacobi_sweeps", cap)
        ctx.eigh(A); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            w, V = ctx.eigh(A)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) / 3 * 1e3
        wn = np.sort(w.cpu().numpy())
        Vn = V.cpu().numpy()
        res = np.linalg.norm(A.cpu().numpy() @ Vn - Vn * w.cpu().numpy()) / np.linalg.norm(wref)
        print(f"d={d} cap={cap:2d} {ms:8.3f} ms  eig relerr={np.max(np.abs(wn - wref)) / wref[-1]:.2e}  resid={res:.2e} "
              f"orth={np.max(np.abs(Vn.T @ Vn - np.eye(d))):.2e}", flush=True)
