"""End-to-end oracle parity at the BASELINE.json configs (B200 only: `-m gpu`), through the legacy C symbols
(pca_float / truncated_svd_float / pca_double) against the reference's own CPU backend run live
(oracle/_ref/libdimreduce4cpu.so, src/cpu_backend.cpp:338-387), at the sizes BASELINE.md section 4 names:

  C2  PCA k=32, 1,000,000 x 256 fp32            FULL size                      (oracle: sgesdd, ~11 s on 16 cores)
  C4  PCA k=32, d=512 fp32                      500,000 rows of the shard      (oracle ~25 s; 2M rows would need ~90 s)
  C3  TSVD k=64, d=1024 fp32                    150,000 rows: 'cusolver' vs oracle-exact at 1e-4; 'power' on a spiked
                                                input (r = 2k) vs oracle-exact at 1e-4 AND the reference's own
                                                35 deg / 1.10x criteria vs the oracle's randomized path
  C5  PCA k=100, d=4096 fp64                    16,384 rows vs the fp64 restatement at 1e-10 (the reference has no
                                                double CPU entry point: "parity unpinned" at fp64 by the reference itself)
  C1  TSVD k=10, 10,000 x 100                   golden (tests/test_gpu_parity.py) + here live with the default algorithm

Tolerances are BASELINE.json's: singular values / explained variance 1e-4 relative (fp32), 1e-10 (fp64); components
entrywise after a common svd_flip where the spectrum is separated, by principal angle otherwise.  Every achieved error
is printed in the assertion message and appended to gpurun_out/parity_report.jsonl."""
from __future__ import annotations

import json
import time
from pathlib import Path

import numpy as np
import pytest

from oracle import cpu_oracle as O

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parents[1]
TOL32, TOL64 = 1e-4, 1e-10


def _report(name: str, **kw) -> None:
    try:
        out = ROOT / "gpurun_out"
        out.mkdir(exist_ok=True)
        with open(out / "parity_report.jsonl", "a") as f:
            f.write(json.dumps(dict(case=name, **{k: (float(v) if isinstance(v, (np.floating, float)) else v) for k, v in kw.items()})) + "\n")
    except OSError:
        pass


def _rel(a, b) -> float:
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / (np.abs(b) + 1e-300)))


def _make(n, d, kind, seed, dtype=np.float32):
    """bench.py's recipe: randn * linspace(1,3,d) (+ linspace(-5,5,d) for PCA)."""
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, d), dtype=np.float32).astype(dtype, copy=False)
    X *= np.linspace(1.0, 3.0, d, dtype=dtype)
    if kind == "pca":
        X += np.linspace(-5.0, 5.0, d, dtype=dtype)
    return X


def _spiked(n, d, k, seed):
    """SURVEY 8d 'spiked': X = randn(n,r) diag(geomspace(10,1,r)) Q^T + 0.1 randn(n,d), r = 2k: a gap after component k
    is not needed for the exact solver, but 'power' (subspace iteration) converges at the rate sigma_{l+1}/sigma_k."""
    rng = np.random.default_rng(seed)
    r = 2 * k
    Qm, _ = np.linalg.qr(rng.standard_normal((d, r)))
    S = np.geomspace(10.0, 1.0, r)
    X = (rng.standard_normal((n, r)) * S) @ Qm.T
    X += 0.1 * rng.standard_normal((n, d))
    return np.ascontiguousarray(X, dtype=np.float32)


def _fit(kind, X, k, alg, **kw):
    import dimreduce4gpu_b200 as pkg

    cls = pkg.PCA if kind == "pca" else pkg.TruncatedSVD
    est = cls(n_components=k, algorithm=alg, n_iter=kw.pop("n_iter", 5), random_state=42, backend="gpu", **kw)
    t0 = time.perf_counter()
    Z = est.fit_transform(X)
    return est, Z, time.perf_counter() - t0


def _oracle(kind, X, k, alg, n_iter=5):
    lib = O.RefLib()
    O.RefLib.set_threads(__import__("os").cpu_count() or 1)
    t0 = time.perf_counter()
    ref = (lib.pca if kind == "pca" else lib.tsvd)(X, k, alg, n_iter, 42)
    return ref, time.perf_counter() - t0


def _compare(name, est, Z, ref, tol, t_gpu, t_cpu, sep_gap=2e-2):
    s, rs = est.singular_values_.astype(np.float64), ref.singular_values.astype(np.float64)
    e_s = _rel(s, rs)
    e_ev = _rel(est.explained_variance_, ref.explained_variance)
    e_evr = _rel(est.explained_variance_ratio_, ref.explained_variance_ratio)
    Cr, Zr = O.svd_flip_v(ref.components, ref.X_transformed)
    Co, Zo = O.svd_flip_v(est.components_, Z)
    gaps = np.abs(np.diff(np.concatenate([rs, [0.0]]))) / rs[0]
    sep = np.minimum(gaps, np.concatenate([[1.0], gaps[:-1]])) > sep_gap
    e_c = float(np.max(np.abs(Cr[sep] - Co[sep]))) if sep.any() else float("nan")
    ang = O.max_principal_angle(Cr, Co)
    if ang >= 1e-2 and len(rs) > 1:   # the cut after component k falls inside a cluster: containment of the first k-1
        ang = O.max_principal_angle(Co, Cr[:-1])
    # scores relative to their own column norm (= sigma), sign-aligned
    e_z = float(np.max(np.linalg.norm(Zr.astype(np.float64) - Zo.astype(np.float64), axis=0) / rs))
    e_m = float(np.max(np.abs(est.mean_ - ref.mean))) if ref.mean is not None else 0.0
    C = est.components_.astype(np.float64)
    e_o = float(np.max(np.abs(C @ C.T - np.eye(C.shape[0]))))
    _report(name, sigma_rel=e_s, ev_rel=e_ev, evr_rel=e_evr, comp_sep_abs=e_c, n_separated=int(sep.sum()), max_angle=ang,
            scores_rel=e_z, mean_abs=e_m, orth=e_o, gpu_s=t_gpu, oracle_s=t_cpu, n=int(Z.shape[0]), d=int(C.shape[1]))
    msg = (f"{name}: sigma_rel={e_s:.2e} ev_rel={e_ev:.2e} evr_rel={e_evr:.2e} comp_abs(sep {int(sep.sum())})={e_c:.2e} "
           f"angle={ang:.2e} scores_rel={e_z:.2e} mean_abs={e_m:.2e} orth={e_o:.2e} (gpu {t_gpu:.2f}s, oracle {t_cpu:.1f}s)")
    print(msg)
    assert e_s < tol, msg
    assert e_ev < 2 * tol and e_evr < 2 * tol, msg
    # flat spectra (relative gaps ~1e-3): individual vectors are conditioned like tol / gap - the subspace is the
    # well-posed object (BASELINE.json: "or by principal-angle bound where eigenvalues are near-degenerate")
    assert ang < 1e-2, msg
    if sep.any():
        assert e_c < 50 * tol, msg
    assert e_o < 50 * tol, msg
    assert e_m < tol * 10, msg
    return msg


def test_config2_full_size_pca_vs_reference_cpu_backend():
    """BASELINE configs[1] at FULL size: 1,000,000 x 256 fp32, PCA k=32."""
    X = _make(1_000_000, 256, "pca", 1234)
    est, Z, tg = _fit("pca", X, 32, "cusolver")
    ref, tc = _oracle("pca", X, 32, "cusolver")
    _compare("C2_full_1Mx256_k32", est, Z, ref, TOL32, tg, tc)


def test_config4_shard_slice_pca_vs_reference_cpu_backend():
    """BASELINE configs[3] family: d=512, k=32, 500,000 rows (the oracle needs ~25 s here; 2M rows ~90 s)."""
    X = _make(500_000, 512, "pca", 1235)
    est, Z, tg = _fit("pca", X, 32, "cusolver")
    ref, tc = _oracle("pca", X, 32, "cusolver")
    _compare("C4_500kx512_k32", est, Z, ref, TOL32, tg, tc)


def test_config3_exact_vs_reference_exact():
    """BASELINE configs[2] family: d=1024, k=64, 'cusolver' on both sides, 150,000 rows."""
    X = _make(150_000, 1024, "tsvd", 1236)
    est, Z, tg = _fit("tsvd", X, 64, "cusolver")
    ref, tc = _oracle("tsvd", X, 64, "cusolver")
    _compare("C3_exact_150kx1024_k64", est, Z, ref, TOL32, tg, tc)


def test_config3_power_on_spiked_input():
    """algorithm='power' at d=1024, k=64 on a spiked input (r = 2k): (a) vs the oracle's EXACT path at 1e-4 (iterated to
    tol), (b) with the default n_iter=5 by the reference's own criteria against its randomized path (subspace angle
    < 35 deg, reconstruction error <= 1.10x: tests/test_sklearn_parity_comprehensive.py:177-210)."""
    n, d, k = 150_000, 1024, 64
    X = _spiked(n, d, k, 1237)
    ex, tc = _oracle("tsvd", X, k, "cusolver")
    est, Z, tg = _fit("tsvd", X, k, "power", n_iter=200, tol=1e-10)
    e_s = _rel(est.singular_values_, ex.singular_values)
    (Cr,), (Co,) = O.svd_flip_v(ex.components), O.svd_flip_v(est.components_)
    ang = O.max_principal_angle(Cr, Co)
    # (b) default iterations vs the oracle's randomized path
    rnd, tr = _oracle("tsvd", X, k, "power", n_iter=5)
    est5, Z5, _ = _fit("tsvd", X, k, "power", n_iter=5)
    (Cn,), (C5,) = O.svd_flip_v(rnd.components), O.svd_flip_v(est5.components_)
    ang5 = O.max_principal_angle(Cn, C5)
    Xd = X.astype(np.float64)
    rec5 = np.linalg.norm(Xd - Z5.astype(np.float64) @ est5.components_.astype(np.float64))
    rec_r = np.linalg.norm(Xd - rnd.X_transformed.astype(np.float64) @ rnd.components.astype(np.float64))
    rec_x = np.linalg.norm(Xd - ex.X_transformed.astype(np.float64) @ ex.components.astype(np.float64))
    _report("C3_power_spiked_150kx1024_k64", sigma_rel_vs_exact=e_s, angle_vs_exact=ang, angle5_vs_randomized=ang5,
            rec5_over_randomized=rec5 / rec_r, rec5_over_exact=rec5 / rec_x, gpu_s=tg, oracle_exact_s=tc, oracle_rand_s=tr)
    msg = (f"C3 power spiked: sigma_rel_vs_exact={e_s:.2e} angle_vs_exact={ang:.2e}; n_iter=5: angle_vs_randomized="
           f"{np.rad2deg(ang5):.2f} deg, recon/randomized={rec5 / rec_r:.4f}, recon/exact={rec5 / rec_x:.4f}")
    print(msg)
    assert e_s < TOL32, msg
    assert ang < 1e-2, msg
    assert ang5 < np.deg2rad(35), msg
    assert rec5 <= 1.10 * rec_r, msg


def test_config1_default_algorithm_live():
    """BASELINE configs[0]: TruncatedSVD n_components=10 with the DEFAULT algorithm ('power') on 10,000 x 100 fp32, exactly
    as a user of the reference calls it.  The oracle takes its exact path here (min(n,d) <= 256, src/cpu_backend.cpp:347)
    and so does the library (csrc/api.cu: same rule), so the default call is held to the full 1e-4 gate."""
    X = np.random.default_rng(1234).standard_normal((10_000, 100)).astype(np.float32) * np.linspace(1, 3, 100, dtype=np.float32)
    ref, tc = _oracle("tsvd", X, 10, "power")
    est, Z, tg = _fit("tsvd", X, 10, "power")
    _compare("C1_10000x100_k10_default_algorithm", est, Z, ref, TOL32, tg, tc)


def test_config5_family_fp64_vs_restatement():
    """BASELINE configs[4] family: PCA k=100, d=4096, fp64, 16,384 rows, through pca_double (DMMA Gram + the d=4096
    eigensolve) vs the fp64 restatement of src/cpu_backend.cpp:359-387 at 1e-10 relative."""
    import dimreduce4gpu_b200 as pkg

    n, d, k = 16_384, 4096, 100
    rng = np.random.default_rng(1238)
    X = rng.standard_normal((n, d)) * np.geomspace(4.0, 0.5, d) + np.linspace(-3.0, 3.0, d)
    t0 = time.perf_counter()
    ref = O.restated_pca(X, k, dtype=np.float64)
    tc = time.perf_counter() - t0
    est = pkg.PCA(n_components=k, algorithm="cusolver", backend="gpu", dtype="float64")
    t0 = time.perf_counter()
    Z = est.fit_transform(X)
    tg = time.perf_counter() - t0
    s, rs = est.singular_values_, ref.singular_values
    e_s = _rel(s, rs)
    e_ev = _rel(est.explained_variance_, ref.explained_variance)
    e_evr = _rel(est.explained_variance_ratio_, ref.explained_variance_ratio)
    e_m = float(np.max(np.abs(est.mean_ - ref.mean)))
    Cr, Zr = O.svd_flip_v(ref.components, ref.X_transformed)
    Co, Zo = O.svd_flip_v(est.components_, Z)
    gaps = np.abs(np.diff(np.concatenate([rs, [0.0]]))) / rs[0]
    sep = np.minimum(gaps, np.concatenate([[1.0], gaps[:-1]])) > 1e-3
    e_c = float(np.max(np.abs(Cr[sep] - Co[sep]))) if sep.any() else float("nan")
    ang = O.max_principal_angle(Cr, Co)
    e_z = float(np.max(np.linalg.norm(Zr - Zo, axis=0) / rs))
    _report("C5_16384x4096_k100_fp64", sigma_rel=e_s, ev_rel=e_ev, evr_rel=e_evr, mean_abs=e_m, comp_sep_abs=e_c,
            n_separated=int(sep.sum()), max_angle=ang, scores_rel=e_z, gpu_s=tg, oracle_s=tc)
    msg = (f"C5 fp64: sigma_rel={e_s:.2e} ev_rel={e_ev:.2e} evr_rel={e_evr:.2e} mean_abs={e_m:.2e} comp_abs(sep {int(sep.sum())})="
           f"{e_c:.2e} angle={ang:.2e} scores_rel={e_z:.2e} (gpu {tg:.2f}s, restatement {tc:.1f}s)")
    print(msg)
    assert e_s < TOL64 and e_ev < 2 * TOL64 and e_evr < 2 * TOL64, msg
    assert e_m < 1e-12, msg
    assert ang < 1e-6, msg
    if sep.any():
        assert e_c < 1e-7, msg


def test_whiten_opt_in_gives_unit_variance_scores_and_roundtrips():
    """PCA(whiten=True): rows of components_ / sqrt(explained_variance_), scores have unit variance, inverse_transform
    undoes it; whiten=False (default) is untouched (the parity oracle ignores the flag)."""
    import dimreduce4gpu_b200 as pkg

    X = _make(20_000, 64, "pca", 9)
    plain = pkg.PCA(n_components=8, backend="gpu")
    Zp = plain.fit_transform(X)
    wh = pkg.PCA(n_components=8, backend="gpu", whiten=True)
    Zw = wh.fit_transform(X)
    np.testing.assert_allclose(wh.singular_values_, plain.singular_values_, rtol=1e-6)
    np.testing.assert_allclose(wh.components_ * np.sqrt(wh.explained_variance_)[:, None], plain.components_, atol=2e-6)
    np.testing.assert_allclose(Zw.std(axis=0, ddof=1), np.ones(8), rtol=2e-4)
    np.testing.assert_allclose(wh.inverse_transform(Zw), plain.inverse_transform(Zp), atol=5e-3)


def test_fit_after_partial_fit_is_not_overwritten_by_stale_chunks():
    import dimreduce4gpu_b200 as pkg

    X1, X2 = _make(5000, 48, "pca", 1), 3.0 * _make(6000, 48, "pca", 2)
    est = pkg.PCA(n_components=4, backend="gpu")
    est.partial_fit(X1)
    est.fit(X2)
    fresh = pkg.PCA(n_components=4, backend="gpu").fit(X2)
    np.testing.assert_allclose(est.singular_values_, fresh.singular_values_, rtol=1e-6)
