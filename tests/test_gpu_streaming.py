"""Streaming / out-of-core fit, B200 only (`-m gpu`): dr4g_partial_fit_* through DeviceContext and
the estimators, and the chunked two-sweep path of the legacy host-pointer symbols (DR4G_STREAM_ROWS), checked
against the reference's CPU backend (live oracle) and against the one-pass fit of the same data.

Tolerances: BASELINE.json's 1e-4 relative (fp32) / 1e-10 (fp64) against the oracle; chunked vs one-pass fit of
the same rows differ only by summation order (1e-5 / 1e-12 relative on singular values)."""
from __future__ import annotations

import os

import numpy as np
import pytest

from oracle import cpu_oracle as O
from test_gpu_parity import TOL32, _check_against, _rel

pytestmark = pytest.mark.gpu


def _data(n, d, seed, dtype=np.float32, offset=True):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d)) * np.linspace(1.0, 3.0, d)
    if offset:
        X += np.linspace(-5.0, 5.0, d)
    return np.ascontiguousarray(X.astype(dtype))


def _ref(kind, X, k):
    if O.REF_SO.is_file():
        return (O.RefLib().pca if kind == "pca" else O.RefLib().tsvd)(X, k, "cusolver")
    return (O.restated_pca if kind == "pca" else O.restated_tsvd)(X, k)


@pytest.mark.parametrize("kind,n,d,k,cuts", [("pca", 20000, 256, 32, (7000, 7001, 19999)), ("tsvd", 9000, 100, 10, (1, 4500)),
                                              ("pca", 3000, 320, 8, (1000, 2000)), ("pca", 600, 33, 5, (599,))])
def test_partial_fit_context_matches_one_pass_fit(kind, n, d, k, cuts):
    """Ragged chunks (a 1-row chunk, a chunk below the pilot size) through dr4g_partial_fit_float."""
    import torch

    import dimreduce4gpu_b200 as pkg

    X = torch.from_numpy(_data(n, d, 5 + n)).cuda()
    center = kind == "pca"
    ctx = pkg.DeviceContext(0)
    whole = ctx.fit(X, k, center=center, algorithm="cusolver")
    ctx.partial_fit_reset()
    edges = (0,) + tuple(cuts) + (n,)
    for a, b in zip(edges[:-1], edges[1:]):
        ctx.partial_fit(X[a:b].contiguous(), center=center)
    assert ctx.partial_fit_rows() == n
    part = ctx.partial_fit_finalize(d, k, center, torch.float32, algorithm="cusolver")
    ctx.sync()
    assert _rel(part["singular_values"].cpu().numpy(), whole["singular_values"].cpu().numpy()) < 1e-5
    assert _rel(part["explained_variance_ratio"].cpu().numpy(), whole["explained_variance_ratio"].cpu().numpy()) < 1e-5
    if center:
        assert float((part["mean"] - whole["mean"]).abs().max()) < 1e-5
    ang = O.max_principal_angle(part["components"].cpu().numpy(), whole["components"].cpu().numpy())
    assert ang < 1e-3
    # the state survives finalize: more chunks may follow
    ctx.partial_fit(X[:100].contiguous(), center=center)
    assert ctx.partial_fit_rows() == n + 100
    ctx.partial_fit_reset()
    assert ctx.partial_fit_rows() == 0


def test_partial_fit_rejects_mismatched_chunks():
    import torch

    import dimreduce4gpu_b200 as pkg

    ctx = pkg.DeviceContext(0)
    ctx.partial_fit_reset()
    ctx.partial_fit(torch.randn(100, 64, device="cuda"), center=True)
    with pytest.raises(RuntimeError, match="does not match"):
        ctx.partial_fit(torch.randn(100, 65, device="cuda"), center=True)
    with pytest.raises(RuntimeError, match="does not match"):
        ctx.partial_fit(torch.randn(100, 64, device="cuda", dtype=torch.float64), center=True)
    ctx.partial_fit_reset()
    with pytest.raises(RuntimeError, match="nothing accumulated"):
        ctx.partial_fit_finalize(64, 4, True, torch.float32)


@pytest.mark.parametrize("kind,n,d,k", [("pca", 12000, 256, 16), ("tsvd", 8000, 130, 7)])
def test_estimator_partial_fit_matches_oracle(kind, n, d, k):
    """numpy chunks through PCA.partial_fit / TruncatedSVD.partial_fit vs the reference CPU backend on all rows."""
    import dimreduce4gpu_b200 as pkg

    X = _data(n, d, 31 + n)
    ref = _ref(kind, X, k)
    cls = pkg.PCA if kind == "pca" else pkg.TruncatedSVD
    est = cls(n_components=k, algorithm="cusolver", random_state=42, backend="gpu")
    for a in range(0, n, 2500):
        est.partial_fit(X[a:a + 2500])
    Z = est.transform(X, center=kind == "pca")   # finalises lazily
    _check_against(est, Z, ref.components, ref.singular_values, ref.explained_variance, ref.explained_variance_ratio,
                   ref.mean, ref.X_transformed, TOL32, n)


def test_estimator_partial_fit_fp64():
    import dimreduce4gpu_b200 as pkg

    n, d, k = 5000, 96, 12
    X = _data(n, d, 77, dtype=np.float64)
    ref = O.restated_pca(X, k, dtype=np.float64)
    est = pkg.PCA(n_components=k, backend="gpu", dtype="float64")
    for a in range(0, n, 1234):
        est.partial_fit(X[a:a + 1234])
    est.finalize_fit()
    assert est.components_.dtype == np.float64
    assert _rel(est.singular_values_, ref.singular_values) < 1e-10
    np.testing.assert_allclose(est.mean_, ref.mean, rtol=0, atol=1e-12)


@pytest.mark.parametrize("kind,n,d,k,rows", [("pca", 20000, 256, 32, 3000), ("tsvd", 10000, 100, 10, 4096), ("pca", 5000, 64, 6, 5000)])
def test_legacy_symbols_out_of_core(kind, n, d, k, rows, monkeypatch):
    """pca_float / truncated_svd_float with DR4G_STREAM_ROWS: two chunked sweeps over the host X (H2D, pass /
    projection and D2H overlapped on three streams) give the reference's outputs, including U and X_transformed."""
    import dimreduce4gpu_b200 as pkg

    X = _data(n, d, 900 + n)
    ref = _ref(kind, X, k)
    monkeypatch.setenv("DR4G_STREAM_ROWS", str(rows))
    cls = pkg.PCA if kind == "pca" else pkg.TruncatedSVD
    est = cls(n_components=k, algorithm="cusolver", random_state=42, backend="gpu")
    Z = est.fit_transform(X)
    monkeypatch.delenv("DR4G_STREAM_ROWS")
    _check_against(est, Z, ref.components, ref.singular_values, ref.explained_variance, ref.explained_variance_ratio,
                   ref.mean, ref.X_transformed, TOL32, n)
    # U = X_transformed / sigma, as the reference's get_tsvd_attr (src/tsvd.cu:171-175)
    np.testing.assert_allclose(est._U * est.singular_values_[None, :], Z, rtol=1e-5, atol=1e-4)
    # and the resident path gives the same scores
    est2 = cls(n_components=k, algorithm="cusolver", random_state=42, backend="gpu")
    Z2 = est2.fit_transform(X)
    assert _rel(est2.singular_values_, est.singular_values_) < 1e-5
    assert float(np.max(np.abs(np.abs(Z2) - np.abs(Z)))) < 1e-3 * float(est.singular_values_[0])
    assert os.environ.get("DR4G_STREAM_ROWS") is None


def test_second_device_in_the_same_process():
    """cudaFuncSetAttribute is per device: a second GPU used by the same process must get its own opt-ins
    (dynamic shared memory above 48 KB, cluster size 16).  Needs two visible GPUs."""
    import torch

    import dimreduce4gpu_b200 as pkg

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    X0 = torch.from_numpy(_data(6000, 256, 3)).cuda(0)
    r0 = pkg.DeviceContext(0).fit(X0, 16, center=True)
    torch.cuda.synchronize(0)
    with torch.cuda.device(1):
        X1 = X0.to("cuda:1")
        c1 = pkg.DeviceContext(1)
        r1 = c1.fit(X1, 16, center=True)
        Z1 = c1.transform(X1, r1["components"], mean=r1["mean"])
        rp = c1.fit(X1, 16, center=False, algorithm="power")
        c1.sync()
    assert Z1.shape == (6000, 16) and rp["singular_values"].shape == (16,)
    assert torch.equal(r0["singular_values"].cpu(), r1["singular_values"].cpu())
    # and through the legacy symbol with gpu_id = 1
    est = pkg.PCA(n_components=8, backend="gpu", gpu_id=1)
    Zh = est.fit_transform(_data(3000, 256, 4))
    assert Zh.shape == (3000, 8) and np.isfinite(Zh).all()


# ---------------------------------------------------------------------------------------------- CSR input
@pytest.mark.parametrize("kind,n,d,k,density", [("tsvd", 20000, 300, 12, 0.02), ("pca", 5000, 96, 8, 0.1), ("tsvd", 700, 33, 5, 0.3)])
def test_csr_input_is_expanded_on_the_device_and_matches_the_dense_reference(kind, n, d, k, density, monkeypatch):
    """scipy CSR in: only the CSR arrays are uploaded, row chunks are expanded by dr4g_csr_to_dense_* and streamed through
    partial_fit; the result equals the reference CPU backend on the densified matrix (which is what the reference does on
    the host, truncated_svd.py:67-70).  Small chunks force several of them."""
    import scipy.sparse as sp

    import dimreduce4gpu_b200 as pkg
    from dimreduce4gpu_b200 import truncated_svd as T
    from oracle import cpu_oracle as O

    rng = np.random.default_rng(n + d)
    Xs = sp.random(n, d, density=density, format="csr", dtype=np.float32, random_state=rng,
                   data_rvs=lambda m: rng.standard_normal(m).astype(np.float32) * 2.0 + 1.0)
    monkeypatch.setattr(T._DeviceCSR, "CHUNK_BYTES", 1 << 20)
    cls = pkg.PCA if kind == "pca" else pkg.TruncatedSVD
    est = cls(n_components=k, algorithm="cusolver", backend="gpu")
    Z = est.fit_transform(Xs)
    Xd = np.asarray(Xs.todense(), dtype=np.float32)
    ref = (O.RefLib().pca if kind == "pca" else O.RefLib().tsvd)(Xd, k, "cusolver") if O.REF_SO.is_file() else \
        (O.restated_pca if kind == "pca" else O.restated_tsvd)(Xd, k)
    rel = float(np.max(np.abs(est.singular_values_ - ref.singular_values) / ref.singular_values))
    assert rel < 1e-4, rel
    Cr, Zr = O.svd_flip_v(ref.components, ref.X_transformed)
    Co, Zo = O.svd_flip_v(est.components_, Z)
    assert O.max_principal_angle(Cr, Co) < 1e-2
    assert np.max(np.abs(np.linalg.norm(Zr, axis=0) - np.linalg.norm(Zo, axis=0)) / ref.singular_values) < 2e-4
    # sparse transform (reference semantics: no centring) equals the dense one
    np.testing.assert_allclose(est.transform(Xs), est.transform(Xd), atol=2e-4 * float(ref.singular_values[0]))


def test_csr_to_dense_kernel_matches_scipy():
    import scipy.sparse as sp
    import torch

    import dimreduce4gpu_b200 as pkg

    rng = np.random.default_rng(0)
    Xs = sp.random(1000, 77, density=0.05, format="csr", dtype=np.float64, random_state=rng)
    ctx = pkg.DeviceContext(0)
    ip = torch.from_numpy(Xs.indptr.astype(np.int64)).cuda()
    ix = torch.from_numpy(Xs.indices.astype(np.int32)).cuda()
    dv = torch.from_numpy(Xs.data).cuda()
    D = ctx.csr_to_dense(ip, ix, dv, 100, 650, 77)
    ctx.sync()
    np.testing.assert_array_equal(D.cpu().numpy(), np.asarray(Xs[100:750].todense()))


def test_stream_switch_orders_arena_reuse_and_threads_share_one_context():
    """the estimators share one cached context per GPU.  (a) A call on another torch stream reuses the
    arena the previous call's kernels may still be reading: dr4g_ctx_set_stream orders the new stream behind the old
    one.  (b) Two Python threads calling through the same context: stream selection + launch is one critical section."""
    import threading

    import torch

    import dimreduce4gpu_b200 as pkg

    ctx = pkg.DeviceContext(0)
    g = torch.Generator(device="cuda").manual_seed(3)
    X1 = torch.randn(300_000, 256, device="cuda", generator=g) * torch.linspace(1, 3, 256, device="cuda")
    X2 = torch.randn(200_000, 512, device="cuda", generator=g) * torch.linspace(1, 2, 512, device="cuda")
    ref1 = {k: v.clone() for k, v in ctx.fit(X1, 16, center=True).items()}
    ref2 = {k: v.clone() for k, v in ctx.fit(X2, 8, center=True).items()}
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(3):
        with torch.cuda.stream(s1):
            o1 = {k: v.clone() for k, v in ctx.fit(X1, 16, center=True).items()}
        with torch.cuda.stream(s2):          # no synchronisation in between: same arena, other stream
            o2 = {k: v.clone() for k, v in ctx.fit(X2, 8, center=True).items()}
        torch.cuda.synchronize()
        assert torch.equal(o1["singular_values"], ref1["singular_values"]) and torch.equal(o1["components"], ref1["components"])
        assert torch.equal(o2["singular_values"], ref2["singular_values"]) and torch.equal(o2["components"], ref2["components"])

    errs = []

    def worker(X, k, ref, stream):
        try:
            with torch.cuda.stream(stream):
                for _ in range(5):
                    o = ctx.fit(X, k, center=True)
                    sv = o["singular_values"].clone()
                    stream.synchronize()
                    if not torch.equal(sv, ref["singular_values"]):
                        errs.append("mismatch")
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))

    t1 = threading.Thread(target=worker, args=(X1, 16, ref1, s1))
    t2 = threading.Thread(target=worker, args=(X2, 8, ref2, s2))
    t1.start(); t2.start(); t1.join(); t2.join()
    assert not errs, errs
