"""K7 on tcgen05 (csrc/project_tc.cu) through dr4g_transform_float against an fp64 product, and against the CUDA-core
kernel it replaces (proj_path=1).  Replaces the reference's SGEMM projection (src/tsvd.cu:152-177,229-260)."""
from __future__ import annotations

import numpy as np
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
    X = torch.randn(n, d, device="cuda", generator=g) * torch.linspace(1, 3, d, device="cuda") + torch.linspace(-5, 5, d, device="cuda")
    Q, _ = torch.linalg.qr(torch.randn(d, k, device="cuda", generator=g, dtype=torch.float64))
    Q = Q.T.contiguous().float()
    mean = X.mean(0).contiguous()
    sigma = torch.linspace(3.0, 1.0, k, device="cuda")
    return X, Q, mean, sigma


# one tile, ragged last tile, fewer rows than a tile, d not a multiple of 64, k = 4 .. 128, more tiles than SMs
SHAPES = [(128, 64, 32), (1, 8, 4), (100, 100, 12), (5000, 256, 32), (70001, 512, 32), (3000, 1024, 64), (2000, 320, 100),
          (999, 132, 128), (40000, 64, 8), (300000, 256, 32),
          # k % 4 != 0 (scalar stores of the score rows): the usual small k of a PCA for plotting, C1's k = 10, k = 50
          (20000, 256, 2), (5000, 100, 10), (7777, 512, 3), (30000, 784, 50), (1000, 64, 1), (2048, 128, 127),
          # k > 128: column slices of the score matrices; d % 4 != 0: zero-padded chunk copies of X (both at once too)
          (20000, 1024, 200), (9000, 512, 129), (20000, 1001, 32), (30000, 50, 10), (5000, 254, 200), (70000, 30, 5)]


@pytest.mark.parametrize("n,d,k", SHAPES)
@pytest.mark.parametrize("centered", [True, False])
def test_project_tc_matches_fp64(n, d, k, centered):
    import torch

    ctx = _ctx(2)   # tensor cores wherever eligible (auto keeps inputs of <= 1 MB on the CUDA-core kernel)
    X, Q, mean, sigma = _problem(n, d, k, 5 * n + d + k)
    Z, U = ctx.transform(X, Q, mean=mean if centered else None, singular_values=sigma, want_u=True)
    ctx.sync()
    assert ctx.last_proj_path == "tcgen05"
    Zr = (X.double() - (mean.double() if centered else 0.0)) @ Q.double().T
    scale = max(float(Zr.abs().max()), 1e-3 * float(X.abs().max()))   # n = 1 centred: Z is exactly 0
    ez = float((Z.double() - Zr).abs().max()) / scale
    eu = float((U.double() - Zr / sigma.double()).abs().max()) / scale
    print(f"n={n} d={d} k={k} centred={centered}: max |Z - Z_fp64| / max|Z| = {ez:.2e}, U {eu:.2e}")
    assert ez < 2e-5 and eu < 2e-5, (ez, eu)


def test_project_tc_agrees_with_cuda_core_kernel_and_zero_sigma():
    import torch

    X, Q, mean, sigma = _problem(20000, 256, 32, 3)
    sigma[5] = 0.0
    a, b = _ctx(0), _ctx(1)
    Za, Ua = a.transform(X, Q, mean=mean, singular_values=sigma, want_u=True)
    Zb, Ub = b.transform(X, Q, mean=mean, singular_values=sigma, want_u=True)
    torch.cuda.synchronize()
    assert a.last_proj_path == "tcgen05" and b.last_proj_path == "cuda_cores"
    assert float((Za - Zb).abs().max()) < 2e-4 * float(Zb.abs().max())
    assert float(Ua[:, 5].abs().max()) == 0.0   # sigma == 0 -> U column 0 (reference src/tsvd.cu:171-175)
    assert float((Ua - Ub).abs().max()) < 2e-4 * float(Ub.abs().max())


def test_ineligible_shapes_fall_back():
    import torch

    ctx = _ctx()
    X, Q, mean, sigma = _problem(500, 50, 6, 1)   # d % 4 != 0: rows of X are not 16-byte aligned
    Z = ctx.transform(X, Q, mean=mean)
    ctx.sync()
    assert ctx.last_proj_path == "cuda_cores"
    np.testing.assert_allclose(Z.cpu().numpy(), ((X - mean) @ Q.T).cpu().numpy(), atol=2e-3)


def test_tiny_inputs_stay_on_cuda_cores_at_fp32_accuracy():
    """proj_path = 0 (auto): n d <= 262144 elements -> plain fp32 FMAs, 1e-6 of max|Z| (the bf16 split gives 2e-6 .. 5e-6)."""
    ctx = _ctx(0)
    X, Q, mean, sigma = _problem(200, 80, 16, 9)
    Z = ctx.transform(X, Q, mean=mean)
    ctx.sync()
    assert ctx.last_proj_path == "cuda_cores"
    Zr = (X.double() - mean.double()) @ Q.double().T
    assert float((Z.double() - Zr).abs().max()) < 1e-6 * float(Zr.abs().max())
