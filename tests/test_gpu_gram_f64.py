"""The fp64 Gram kernel (csrc/gram_f64.cu: 128 x 128 DMMA tiles, cp.async ring, persistent schedule with row-split tail
units reduced in a fixed order) through dr4g_gram_double / dr4g_fit_double against an fp64 product of the same inputs,
and against the first-generation 64 x 64 kernel (gram_path=5).  Replaces cublasDgemm of the reference's double entry points
(src/matrix.cu:126 via src/tsvd.cu:324)."""
from __future__ import annotations

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ctx(path=0):
    import dimreduce4gpu_b200 as pkg

    ctx = pkg.DeviceContext(0)
    if path:
        ctx.set_option("gram_path", path)
    return ctx


def _make(n, d, seed, offset=0.5):
    import torch

    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(n, d, device="cuda", generator=g, dtype=torch.float64) * torch.linspace(1, 3, d, device="cuda", dtype=torch.float64) + offset


# one tile; ragged d (partial last block, d % 128 != 0); fewer rows than a stage; rows not a multiple of 16; more tile
# pairs than SMs (d = 2304: 171 pairs = one full round + row-split tail); d = 4096 (BASELINE config 5: 528 pairs)
SHAPES = [(1, 2), (5, 6), (100, 128), (4099, 100), (3001, 130), (777, 256), (20000, 384), (9000, 1030), (3000, 2304), (2500, 4096)]


@pytest.mark.parametrize("n,d", SHAPES)
@pytest.mark.parametrize("shifted", [False, True])
def test_gram_f64_matches_fp64_product(n, d, shifted):
    import torch

    ctx = _ctx()
    X = _make(n, d, 11 * n + d)
    shift = X[: min(n, 4096)].mean(0).contiguous() if shifted else None
    G = ctx.gram(X, shift)
    ctx.sync()
    assert ctx.last_gram_path == "dmma_f64_128"
    Xd = X - (shift if shifted else 0.0)
    Gr = Xd.T @ Xd
    dg = torch.sqrt(torch.diag(Gr))
    den = (dg[:, None] * dg[None, :]).clamp_min(1e-30 + 1e-13 * float((X * X).sum()))   # n = 1 shifted: Gr is exactly 0
    err = float(((G - Gr).abs() / den).max())
    assert err < 5e-14, err
    assert float((G - G.T).abs().max()) == 0.0


def test_gram_f64_is_deterministic_and_matches_the_first_kernel():
    import torch

    X = _make(50000, 640, 3)
    shift = X[:4096].mean(0).contiguous()
    a = _ctx().gram(X, shift).clone()
    b = _ctx().gram(X, shift).clone()
    old = _ctx(5).gram(X, shift).clone()
    torch.cuda.synchronize()
    assert torch.equal(a, b)   # fixed summation order: no atomics
    dg = torch.sqrt(torch.diag(old))
    assert float(((a - old).abs() / (dg[:, None] * dg[None, :])).max()) < 5e-14


def test_fit_double_uses_the_new_kernel_and_odd_d_falls_back():
    import torch

    ctx = _ctx()
    X = _make(6000, 192, 5)
    ctx.fit(X, 8, center=True)
    ctx.sync()
    assert ctx.last_gram_path == "dmma_f64_128"
    Xo = _make(3000, 37, 6)    # odd d: rows are not 16-byte aligned -> first-generation kernel
    r = ctx.fit(Xo, 5, center=True)
    ctx.sync()
    assert ctx.last_gram_path == "generic_dmma"
    s = torch.linalg.svdvals(Xo - Xo.mean(0))[:5]
    np.testing.assert_allclose(r["singular_values"].cpu().numpy(), s.cpu().numpy(), rtol=1e-11)
    Xp = _make(3000, 333, 7)   # odd d > 128: zero-padded chunk copies through the 128 x 128 DMMA kernel
    r = ctx.fit(Xp, 5, center=True)
    ctx.sync()
    assert ctx.last_gram_path == "dmma_f64_128"
    s = torch.linalg.svdvals(Xp - Xp.mean(0))[:5]
    np.testing.assert_allclose(r["singular_values"].cpu().numpy(), s.cpu().numpy(), rtol=1e-11)
