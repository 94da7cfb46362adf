"""GPU probe: tcgen05 Gram kernel vs an fp64 product, per shape, with timing.  Usage:
    python tools/gram_probe.py [--shapes n,d;n,d...] [--flush rows]"""
import argparse
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import dimreduce4gpu_b200 as pkg  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="64,128;200,128;4096,128;5000,256;4099,100;3000,320;8192,512;4096,1024;1000000,256")
    ap.add_argument("--flush", type=int, default=0)
    ap.add_argument("--path", type=int, default=2)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    ctx = pkg.DeviceContext(0)
    ctx.set_option("gram_path", a.path)
    if a.flush:
        ctx.set_option("flush_rows", a.flush)
    for spec in a.shapes.split(";"):
        n, d = map(int, spec.split(","))
        g = torch.Generator(device="cuda").manual_seed(n * 7 + d)
        X = torch.randn(n, d, device="cuda", generator=g) * torch.linspace(1, 3, d, device="cuda") + 0.5
        for use_shift in (False, True):
            shift = X[: min(n, 4096)].mean(0).contiguous() if use_shift else None
            G = ctx.gram(X, shift)
            ctx.sync()
            Xd = X.double() - (shift.double() if use_shift else 0.0)
            Gr = Xd.T @ Xd
            dg = torch.sqrt(torch.diag(Gr))
            err = ((G - Gr).abs() / (dg[:, None] * dg[None, :])).max().item()
            sym = (G - G.T).abs().max().item()
            if a.verbose:
                print("G  ", G[:2, :4].tolist(), G[d - 1, d - 4:].tolist())
                print("Gr ", Gr[:2, :4].tolist(), Gr[d - 1, d - 4:].tolist())
                nz = (G != 0).sum().item()
                print("nonzero", nz, "of", d * d, "nan", torch.isnan(G).sum().item())
                if use_shift:
                    res = ctx.fit(X, min(4, d), center=True)
                    ctx.sync()
                    print("mean err", (res["mean"].double() - X.double().mean(0)).abs().max().item(),
                          "sv", res["singular_values"].tolist())
            t = []
            for _ in range(a.reps):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                ctx.gram(X, shift)
                ctx.sync()
                t.append(time.perf_counter() - t0)
            ms = sorted(t)[len(t) // 2] * 1e3
            print(f"n={n:8d} d={d:5d} shift={int(use_shift)} path={ctx.last_gram_path:12s} rel_err={err:.3e} asym={sym:.2e} "
                  f"wall_ms={ms:8.3f} rows/s={n / ms * 1e3:.3e} GB/s={n * d * 4 / ms / 1e6:8.1f}", flush=True)


if __name__ == "__main__":
    main()
