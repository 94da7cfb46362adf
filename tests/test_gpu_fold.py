"""Row folding of narrow fp32 inputs (csrc/api.cu: gram_fold_factor; csrc/finalize.cu: unfold_moments_kernel).

For d <= 128 and n d >= 2^24 the streamed pass reads f consecutive rows as one row of f d <= 256 columns and takes
G, s, q from the diagonal blocks of the folded moments.  Checked here: the fold factor that was used, the folded fit
against the same fit with folding switched off, and against the reference CPU backend (`oracle/_ref`) on the same rows -
PCA and TruncatedSVD, d % 4 == 0 and not, row counts that leave a tail, chunked partial_fit, the legacy host symbols."""
from __future__ import annotations

import numpy as np
import pytest

from oracle import cpu_oracle as O

pytestmark = pytest.mark.gpu

TOL = 1e-4   # BASELINE.json north_star: singular values / explained variance, fp32


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / (np.abs(b) + 1e-300)))


def _data(n, d, seed, offset=True):
    import torch

    g = torch.Generator(device="cuda").manual_seed(seed)
    X = torch.randn(n, d, device="cuda", generator=g) * torch.linspace(1.0, 3.0, d, device="cuda")
    if offset:
        X = X + torch.linspace(-5.0, 5.0, d, device="cuda")
    return X.contiguous()


def _np(r):
    return {k: (v.double().cpu().numpy() if v is not None else None) for k, v in r.items()}


# (n, d, k, expected fold factor): n d >= 2^24, tails of 0 .. f-1 rows
SHAPES = [(131072 + 1, 128, 16, 2), (524288 + 5, 32, 8, 8), (2097152 + 17, 8, 4, 32), (170000, 100, 10, 2),
          (350003, 48, 6, 5), (2800001, 6, 3, 42), (5600003, 3, 2, 84), (16777216 + 100, 1, 1, 256), (140000, 127, 5, 0)]


@pytest.mark.parametrize("center", [True, False], ids=["pca", "tsvd"])
@pytest.mark.parametrize("n,d,k,f", SHAPES)
def test_folded_fit_matches_unfolded_fit_and_reference(n, d, k, f, center):
    import dimreduce4gpu_b200 as pkg

    X = _data(n, d, 7 * d + k, offset=center)
    ctx = pkg.DeviceContext(0)
    a = _np(ctx.fit(X, k, center=center, algorithm="cusolver"))
    ctx.sync()
    assert ctx.last_fold == f, f"fold factor {ctx.last_fold}, expected {f}"
    if f:
        assert ctx.last_gram_path == "tcgen05_pair_bf16x3"
    ctx.set_option("fold", 0)
    b = _np(ctx.fit(X, k, center=center, algorithm="cusolver"))
    ctx.sync()
    assert ctx.last_fold == 0
    ctx.set_option("fold", 1)
    e_s = _rel(a["singular_values"], b["singular_values"])
    e_evr = _rel(a["explained_variance_ratio"], b["explained_variance_ratio"])
    msg = f"n={n} d={d} k={k} f={f}: sigma {e_s:.2e}, evr {e_evr:.2e} (folded vs unfolded)"
    assert e_s < 3e-5 and e_evr < 6e-5, msg
    if center:
        assert np.max(np.abs(a["mean"] - b["mean"])) < 1e-5 * (1 + np.abs(b["mean"]).max())
    # the reference CPU backend on the first rows of the same matrix (it finishes these in seconds); sigma scales with sqrt(n)
    m = min(n, 400000)
    Xh = X[:m].cpu().numpy()
    ref = O.RefLib().pca(Xh, k, "cusolver") if center else O.RefLib().tsvd(Xh, k, "cusolver")
    c = _np(ctx.fit(X[:m], k, center=center, algorithm="cusolver"))
    ctx.sync()
    e_ref = _rel(c["singular_values"], ref.singular_values)
    assert e_ref < TOL, f"{msg}; sigma vs reference backend on {m} rows {e_ref:.2e}"
    if d > 1 and k < d:
        ang = O.max_principal_angle(a["components"], b["components"])
        assert ang < 5e-3, f"{msg}; principal angle folded vs unfolded {ang:.2e}"


def test_folded_fit_through_the_reference_backend_at_full_size():
    """The whole folded matrix against `oracle/_ref` (2M x 16 runs through sgesdd in a few seconds)."""
    import dimreduce4gpu_b200 as pkg

    n, d, k = 2097152 + 3, 16, 5
    X = _data(n, d, 99)
    ctx = pkg.DeviceContext(0)
    a = _np(ctx.fit(X, k, center=True, algorithm="cusolver"))
    ctx.sync()
    assert ctx.last_fold == 16
    ref = O.RefLib().pca(X.cpu().numpy(), k, "cusolver")
    e_s, e_ev = _rel(a["singular_values"], ref.singular_values), _rel(a["explained_variance"], ref.explained_variance)
    e_evr = _rel(a["explained_variance_ratio"], ref.explained_variance_ratio)
    ang = O.max_principal_angle(a["components"], ref.components)
    msg = f"sigma {e_s:.2e} ev {e_ev:.2e} evr {e_evr:.2e} angle {ang:.2e}"
    assert e_s < TOL and e_ev < 2 * TOL and e_evr < 2 * TOL and ang < 1e-2, msg
    assert np.max(np.abs(a["mean"] - ref.mean)) < TOL * (1e-2 + np.abs(ref.mean).max()), msg


def test_folded_chunks_in_partial_fit_match_the_one_pass_fit():
    import dimreduce4gpu_b200 as pkg

    n, d, k = 1500007, 24, 6
    X = _data(n, d, 5)
    ctx = pkg.DeviceContext(0)
    one = _np(ctx.fit(X, k, center=True, algorithm="cusolver"))
    ctx.sync()
    assert ctx.last_fold == 10
    ctx.partial_fit_reset()
    for lo, hi in [(0, 700001), (700001, 700900), (700900, n)]:   # folded, unfolded (small), folded
        ctx.partial_fit(X[lo:hi], center=True)
    two = _np(ctx.partial_fit_finalize(d, k, center=True, dtype=X.dtype))
    ctx.sync()
    assert _rel(two["singular_values"], one["singular_values"]) < 2e-5
    assert np.max(np.abs(two["mean"] - one["mean"])) < 1e-5
    assert O.max_principal_angle(two["components"], one["components"]) < 5e-3


def test_legacy_symbol_folds_its_chunks():
    """`PCA(...).fit_transform(np.ndarray)` (pca_float, host pointers): the 128 MB device chunks of a narrow matrix are
    folded; scores and model against the reference backend."""
    import dimreduce4gpu_b200 as pkg

    rng = np.random.default_rng(3)
    n, d, k = 1200011, 32, 8
    X = (rng.standard_normal((n, d), dtype=np.float32) * np.linspace(1, 3, d, dtype=np.float32) + np.linspace(-2, 2, d, dtype=np.float32))
    est = pkg.PCA(n_components=k, backend="gpu")
    Z = est.fit_transform(X)
    ref = O.RefLib().pca(X, k, "cusolver")
    assert _rel(est.singular_values_, ref.singular_values) < TOL
    assert _rel(est.explained_variance_ratio_, ref.explained_variance_ratio) < 2 * TOL
    Cr, Zr = O.svd_flip_v(ref.components, ref.X_transformed)
    Co, Zo = O.svd_flip_v(est.components_, Z)
    assert O.max_principal_angle(Cr, Co) < 1e-2
    assert np.max(np.abs(Zr - Zo)) < 5e-2 * ref.singular_values[0] / np.sqrt(n)   # 5 % of a typical leading score


@pytest.mark.parametrize("n,d,k,f", [(2097152 + 5, 8, 2, 32), (524288 + 3, 32, 8, 8), (700001, 24, 5, 10), (5600003, 3, 2, 64),
                                     (16777216 + 9, 1, 1, 128), (1048576 + 1, 16, 16, 8), (300000, 64, 8, 0)])
@pytest.mark.parametrize("centered", [True, False])
def test_folded_projection_matches_fp64_and_the_unfolded_kernel(n, d, k, f, centered):
    """K7 on narrow inputs (d <= 32): f rows against the block-diagonal image of the components; Z and U land where the
    unfolded kernel puts them."""
    import torch

    import dimreduce4gpu_b200 as pkg

    g = torch.Generator(device="cuda").manual_seed(n + d + k)
    X = _data(n, d, 3 * n + d, offset=True)
    Q, _ = torch.linalg.qr(torch.randn(d, k, device="cuda", generator=g, dtype=torch.float64))
    Q = Q.T.contiguous().float()
    mean = X[:100000].mean(0).contiguous() if centered else None
    sigma = torch.linspace(3.0, 1.0, k, device="cuda")
    ctx = pkg.DeviceContext(0)
    Z, U = ctx.transform(X, Q, mean=mean, singular_values=sigma, want_u=True)
    ctx.sync()
    assert ctx.last_fold == f, f"fold factor {ctx.last_fold}, expected {f}"
    assert ctx.last_proj_path == "tcgen05"
    ctx.set_option("fold", 0)
    Z0, U0 = ctx.transform(X, Q, mean=mean, singular_values=sigma, want_u=True)
    ctx.sync()
    m = min(n, 300000)
    rows = torch.cat([torch.arange(m, device="cuda"), torch.arange(n - 1000, n, device="cuda")])   # head and the tail rows
    Zr = (X[rows].double() - (mean.double() if centered else 0.0)) @ Q.double().T
    scale = float(Zr.abs().max())
    ez = float((Z[rows].double() - Zr).abs().max()) / scale
    eu = float((U[rows].double() - Zr / sigma.double()).abs().max()) / scale
    e0 = float((Z - Z0).abs().max()) / scale
    msg = f"n={n} d={d} k={k} f={f}: Z {ez:.2e}, U {eu:.2e} vs fp64; folded vs unfolded {e0:.2e}"
    assert ez < 2e-5 and eu < 2e-5 and e0 < 2e-5, msg
    assert float((U - U0).abs().max()) / scale < 2e-5, msg


@pytest.mark.parametrize("center", [True, False], ids=["pca", "tsvd"])
@pytest.mark.parametrize("n,d,k,f", [(131072 + 3, 64, 8, 4), (300007, 32, 6, 8), (600011, 15, 4, 16), (70000, 128, 12, 2),
                                     (200001, 45, 5, 4), (8388608 + 5, 1, 1, 256), (60000, 129, 5, 0)])
def test_folded_fp64_fit_matches_unfolded_fit_and_fp64_eigenvalues(n, d, k, f, center):
    """fp64 inputs with d <= 128 and n d >= 2^23: f rows as one row of f d <= 256 columns on the 128 x 128 DMMA tiles
    (csrc/api.cu: gram_fold_factor for double); gates of the fp64 path: 1e-10 on sigma (BASELINE.json north_star)."""
    import torch

    import dimreduce4gpu_b200 as pkg

    X = _data(n, d, 11 * d + k, offset=center).double()
    X = (X + 1e-3 * torch.randn(n, d, device="cuda", dtype=torch.float64, generator=torch.Generator(device="cuda").manual_seed(n))).contiguous()
    ctx = pkg.DeviceContext(0)
    a = _np(ctx.fit(X, k, center=center, algorithm="cusolver"))
    ctx.sync()
    assert ctx.last_fold == f, f"fold factor {ctx.last_fold}, expected {f}"
    ctx.set_option("fold", 0)
    b = _np(ctx.fit(X, k, center=center, algorithm="cusolver"))
    ctx.sync()
    assert ctx.last_fold == 0
    mu = X.mean(0) if center else torch.zeros(d, device="cuda", dtype=torch.float64)
    ev = torch.linalg.eigvalsh((X - mu).T @ (X - mu)).flip(0)[:k].clamp(min=0).sqrt().cpu().numpy()
    e_ab = _rel(a["singular_values"], b["singular_values"])
    e_ref = _rel(a["singular_values"], ev)
    msg = f"n={n} d={d} k={k} f={f}: sigma folded vs unfolded {e_ab:.1e}, vs fp64 eigenvalues {e_ref:.1e}"
    assert e_ab < 1e-11 and e_ref < 1e-10, msg
    assert _rel(a["explained_variance_ratio"], b["explained_variance_ratio"]) < 1e-11, msg
    if center:
        assert np.max(np.abs(a["mean"] - mu.cpu().numpy())) < 1e-12 * (1 + float(mu.abs().max())), msg
    if d > 1 and k < d:
        assert O.max_principal_angle(a["components"], b["components"]) < 1e-8, msg
