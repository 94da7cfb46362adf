"""K7 for fp64 input on the DMMA pipe (csrc/project_f64.cu) through dr4g_transform_double against a torch fp64 product
of the same inputs, and against the CUDA-core kernel it replaces (proj_path=1).  Replaces the reference's DGEMM
projection of the double entry points (src/tsvd.cu:152-177,229-260, src/matrix.cu:126).  Tolerance: 1e-12 of max |Z|
(BASELINE.json asks 1e-10 for fp64 results)."""
from __future__ import annotations

import pytest

pytestmark = pytest.mark.gpu


def _ctx(path=0):
    import dimreduce4gpu_b200 as pkg

    ctx = pkg.DeviceContext(0)
    ctx.set_option("proj_path", path)
    return ctx


def _problem(n, d, k, seed):
    import torch

    g = torch.Generator(device="cuda").manual_seed(seed)
    dt = torch.float64
    X = torch.randn(n, d, device="cuda", generator=g, dtype=dt) * torch.linspace(1, 3, d, device="cuda", dtype=dt) \
        + torch.linspace(-5, 5, d, device="cuda", dtype=dt)
    Q, _ = torch.linalg.qr(torch.randn(d, k, device="cuda", generator=g, dtype=dt))
    Q = Q.T.contiguous()
    mean = X.mean(0).contiguous()
    sigma = torch.linspace(3.0, 1.0, k, device="cuda", dtype=dt)
    return X, Q, mean, sigma


# one tile, a single row, ragged last tile, d not a multiple of 16, every accumulator-tile count class (k = 1 .. 128),
# config 5's shape (d = 4096, k = 100), more tiles than SMs
SHAPES = [(128, 64, 32), (1, 8, 4), (100, 100, 12), (5000, 256, 32), (3001, 4096, 100), (2000, 322, 7), (999, 130, 128),
          (40000, 64, 1), (30000, 512, 64), (777, 1000, 57),
          # k > 128: column slices; odd d: zero-padded chunk copies
          (4000, 256, 130), (3001, 333, 100), (20000, 99, 10), (2500, 1001, 200)]


@pytest.mark.parametrize("n,d,k", SHAPES)
@pytest.mark.parametrize("centered", [True, False])
def test_project_f64_matches_torch_fp64(n, d, k, centered):
    ctx = _ctx()
    X, Q, mean, sigma = _problem(n, d, k, 5 * n + d + k)
    Z, U = ctx.transform(X, Q, mean=mean if centered else None, singular_values=sigma, want_u=True)
    ctx.sync()
    assert ctx.last_proj_path == "dmma_f64"
    Zr = (X - (mean if centered else 0.0)) @ Q.T
    scale = max(float(Zr.abs().max()), 1e-3 * float(X.abs().max()))
    ez = float((Z - Zr).abs().max()) / scale
    eu = float((U - Zr / sigma).abs().max()) / scale
    print(f"n={n} d={d} k={k} centred={centered}: max |Z - Z_ref| / max|Z| = {ez:.2e}, U {eu:.2e}")
    assert ez < 1e-12 and eu < 1e-12, (ez, eu)


def test_project_f64_agrees_with_cuda_core_kernel_and_zero_sigma():
    import torch

    X, Q, mean, sigma = _problem(20000, 256, 40, 3)
    sigma[5] = 0.0
    a, b = _ctx(0), _ctx(1)
    Za, Ua = a.transform(X, Q, mean=mean, singular_values=sigma, want_u=True)
    Zb, Ub = b.transform(X, Q, mean=mean, singular_values=sigma, want_u=True)
    torch.cuda.synchronize()
    assert a.last_proj_path == "dmma_f64" and b.last_proj_path == "cuda_cores"
    assert float((Za - Zb).abs().max()) < 1e-12 * float(Zb.abs().max())
    assert float(Ua[:, 5].abs().max()) == 0.0   # sigma == 0 -> U column 0 (reference src/tsvd.cu:171-175)
    assert float((Ua - Ub).abs().max()) < 1e-12 * float(Ub.abs().max())


def test_odd_d_falls_back():
    ctx = _ctx()
    X, Q, mean, sigma = _problem(500, 49, 6, 1)
    Z = ctx.transform(X, Q, mean=mean)
    ctx.sync()
    assert ctx.last_proj_path == "cuda_cores"
    assert float((Z - (X - mean) @ Q.T).abs().max()) < 1e-10
