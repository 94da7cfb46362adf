"""The tcgen05 Gram + column-statistics kernels (csrc/gram_tc2.cu: CTA pairs, d > 128 and d % 32 == 0;
csrc/gram_tc.cu: single CTA, the other shapes) through the C ABI (dr4g_gram_float, dr4g_fit_float
with gram_path=2 / 3) against an fp64 product of the same inputs.

Tolerance: 3xTF32 with 1024-row flush windows reproduces X^T X to about 2e-5 of sqrt(G_ii G_jj)
(the tensor core truncates its fp32 accumulation, about 1.6e-8 relative per accumulated row); the
gate is 3e-5, inside what the 1e-4 singular-value tolerance of BASELINE.json needs (sigma^2 are
eigenvalues of G, so sigma moves by half of that).  With flush_rows=128 the error is ~3e-6."""
from __future__ import annotations

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GRAM_TOL = 3e-5


def _ctx(path=2, **opts):
    import dimreduce4gpu_b200 as pkg

    ctx = pkg.DeviceContext(0)
    ctx.set_option("gram_path", path)
    for k, v in opts.items():
        ctx.set_option(k, v)
    return ctx


def _make(n, d, seed, offset=0.5):
    import torch

    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(n, d, device="cuda", generator=g) * torch.linspace(1, 3, d, device="cuda") + offset


def _rel_err(G, Xd):
    import torch

    Gr = Xd.T @ Xd
    dg = torch.sqrt(torch.diag(Gr))
    return float(((G - Gr).abs() / (dg[:, None] * dg[None, :]).clamp_min(1e-30)).max())


# shapes: one/two/three/four/eight column blocks, ragged d (d % 128 != 0, d % 32 != 0), ragged n
# (n % 64 != 0, n < one chunk, n < number of CTAs * chunk), and a long stream with many flush windows
# pair kernel: fewer rows than one stage, odd number of 256-column super-blocks (768), a partial last block
# (320, 384), several flush windows of the diagonal jobs (600000 x 256: 8108 rows per cluster, window 4096)
SHAPES = [(1, 4), (3, 8), (64, 128), (200, 128), (4099, 100), (5000, 256), (3001, 320), (2500, 384),
          (8192, 512), (3000, 516), (4096, 1024), (300000, 256), (70000, 512),
          (10, 256), (33, 512), (1000, 768), (600000, 256), (20000, 2048),
          # ragged d through the pair kernel (d % 32 != 0: the last row goes through gram_tail_rows_kernel): MNIST's 784,
          # 1000, a partial last 128-block with a partial last group, two rows only
          (30000, 784), (12345, 1000), (5000, 132), (2, 516), (70000, 300)]


@pytest.mark.parametrize("n,d", SHAPES)
@pytest.mark.parametrize("shifted", [False, True])
def test_gram_tc_matches_fp64(n, d, shifted):
    ctx = _ctx()
    X = _make(n, d, 11 * n + d)
    shift = X[: min(n, 4096)].mean(0).contiguous() if shifted else None
    G = ctx.gram(X, shift)
    ctx.sync()
    assert ctx.last_gram_path.startswith("tcgen05")
    Xd = X.double() - (shift.double() if shifted else 0.0)
    assert _rel_err(G, Xd) < GRAM_TOL
    assert float((G - G.T).abs().max()) == 0.0  # mirrored, exactly symmetric


@pytest.mark.parametrize("n,d", [(5000, 256), (3001, 320), (9000, 512), (4096, 1024)])
def test_pair_kernel_agrees_with_single_cta_kernel(n, d):
    """gram_path=2 takes the CTA-pair kernel for these shapes, gram_path=3 forces the single-CTA one."""
    import torch

    X = _make(n, d, 17 * n + d)
    cp, cs = _ctx(2), _ctx(3)
    Gp = cp.gram(X).clone()
    Gs = cs.gram(X).clone()
    torch.cuda.synchronize()
    assert cp.last_gram_path.startswith("tcgen05_pair") and cs.last_gram_path == "tcgen05"
    dg = torch.sqrt(torch.diag(Gs))
    assert float(((Gp - Gs).abs() / (dg[:, None] * dg[None, :])).max()) < GRAM_TOL


@pytest.mark.parametrize("n,d", [(5000, 256), (3001, 320), (70000, 512), (4096, 1024), (33, 512), (300000, 256)])
@pytest.mark.parametrize("split,name,tol", [(1, "tcgen05_pair", GRAM_TOL), (2, "tcgen05_pair_bf16x3", GRAM_TOL)])
def test_pair_kernel_both_operand_splits(n, d, split, name, tol):
    """gram_split=1: 3xTF32 on kind::tf32; gram_split=2 (the default): two bf16 terms, three kind::f16 products.  Both
    against the fp64 product at the same gate; the achieved error is printed."""
    ctx = _ctx(2, gram_split=split)
    X = _make(n, d, 13 * n + d)
    shift = X[: min(n, 4096)].mean(0).contiguous()
    G = ctx.gram(X, shift)
    ctx.sync()
    assert ctx.last_gram_path == name
    err = _rel_err(G, X.double() - shift.double())
    print(f"n={n} d={d} {name}: max |G - G_fp64| / sqrt(G_ii G_jj) = {err:.2e}")
    assert err < tol, err
    assert float((G - G.T).abs().max()) == 0.0


def test_bf16x3_split_on_badly_scaled_columns():
    """Column scales from 1e-6 to 1e6 and a duplicated column: bf16 keeps the fp32 exponent range, so the error stays
    relative to sqrt(G_ii G_jj) per entry; the duplicate must give an exactly symmetric, consistent 2 x 2 block."""
    import torch

    n, d = 40000, 256
    X = _make(n, d, 77, offset=0.0) * torch.logspace(-6, 6, d, device="cuda")
    X[:, 17] = X[:, 200]
    ctx = _ctx(2, gram_split=2)
    G = ctx.gram(X)
    ctx.sync()
    assert ctx.last_gram_path == "tcgen05_pair_bf16x3"
    err = _rel_err(G, X.double())
    assert err < GRAM_TOL, err
    a, b, c = float(G[17, 17]), float(G[200, 200]), float(G[17, 200])
    assert abs(a - b) <= 1e-7 * a and abs(a - c) <= 1e-7 * a, (a, b, c)


def test_gram_tc_agrees_with_generic_dmma_path():
    X = _make(20000, 384, 5)
    Gt = _ctx(2).gram(X)
    Gg = _ctx(1).gram(X)
    import torch

    torch.cuda.synchronize()
    dg = torch.sqrt(torch.diag(Gg))
    assert float(((Gt - Gg).abs() / (dg[:, None] * dg[None, :])).max()) < GRAM_TOL


@pytest.mark.parametrize("n,d", [(5000, 256), (4099, 100), (9000, 640)])
def test_fused_column_statistics(n, d):
    """mean_ and the total variance come from the same pass (sums / sums of squares of x - shift)."""
    ctx = _ctx()
    X = _make(n, d, 3, offset=7.0)
    res = ctx.fit(X, min(8, d), center=True)
    ctx.sync()
    assert ctx.last_gram_path.startswith("tcgen05")
    Xd = X.double()
    assert float((res["mean"].double() - Xd.mean(0)).abs().max()) < 2e-6
    tot = float(Xd.var(0, unbiased=True).sum())
    evr = res["explained_variance_ratio"].double()
    ev = res["explained_variance"].double()
    np.testing.assert_allclose(float((ev / evr).mean()), tot, rtol=2e-5)


def test_flush_window_bounds_the_accumulation_bias():
    """The tensor core truncates its fp32 accumulation: the error grows with the flush window,
    which is why the accumulators are flushed every flush_rows rows."""
    X = _make(400000, 256, 9)
    Xd = X.double()
    errs = {}
    for rows in (128, 8192):
        ctx = _ctx(2, flush_rows=rows)
        G = ctx.gram(X)
        ctx.sync()
        errs[rows] = _rel_err(G, Xd)
    assert errs[128] < 5e-6
    assert errs[128] < errs[8192]


def test_full_size_config2_linearity_and_trace():
    """BASELINE configs[1] size (1M x 256): properties that need no oracle - the Gram of the stacked
    halves is the sum of the halves' Grams, and trace(G) = sum of squares."""
    import torch

    ctx = _ctx()
    X = _make(1_000_000, 256, 21)
    G = ctx.gram(X).clone()
    G1 = ctx.gram(X[:500_000]).clone()
    G2 = ctx.gram(X[500_000:]).clone()
    ctx.sync()
    dg = torch.sqrt(torch.diag(G))
    assert float(((G1 + G2 - G).abs() / (dg[:, None] * dg[None, :])).max()) < GRAM_TOL
    ssq = float((X.double() ** 2).sum())
    assert abs(float(torch.trace(G)) - ssq) / ssq < GRAM_TOL
