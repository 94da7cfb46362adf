"""Deterministic input generators shared by tests/golden/make_golden.py (run against the
reference in the build container) and the parity tests (run anywhere).  numpy's PCG64 +
standard_normal stream is stable across numpy versions; every golden file also records the
sha256 of the input it was generated from so a drift would be caught, not silently compared."""
from __future__ import annotations

import hashlib

import numpy as np


def comprehensive_matrix(n, m, k, rank_deficient, dtype, seed):
    """Same construction as the reference's _make_matrix
    (tests/test_sklearn_parity_comprehensive.py:48-62)."""
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((n, m)).astype(dtype, copy=False)
    X *= np.linspace(1.0, 3.0, m, dtype=dtype)
    if rank_deficient:
        A = rng.standard_normal((m, k)).astype(dtype, copy=False)
        X = (X @ A) @ A.T
    return X


def config1_matrix(seed=1234):
    """BASELINE.json configs[0]: 10,000 x 100 fp32 dense random matrix."""
    rng = np.random.default_rng(seed)
    return rng.standard_normal((10000, 100)).astype(np.float32)


def offset_matrix(seed=5):
    """PCA centring stress: |mean| up to 50 sigma (SURVEY 7.3 hard part 5)."""
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((2000, 96)).astype(np.float32) * np.linspace(1.0, 3.0, 96, dtype=np.float32)
    return (X + np.linspace(-50.0, 50.0, 96, dtype=np.float32)).astype(np.float32)


def spiked_matrix(n=3000, d=320, r=24, seed=11):
    """Gapped spectrum (SURVEY 8d 'spiked'): low-rank signal + small noise, d > 256 so that
    algorithm='power' takes the reference's randomized path (src/cpu_backend.cpp:347)."""
    rng = np.random.default_rng(seed)
    Qm, _ = np.linalg.qr(rng.standard_normal((d, r)))
    S = np.geomspace(10.0, 1.0, r)
    X = rng.standard_normal((n, r)) * S @ Qm.T + 0.05 * rng.standard_normal((n, d))
    return X.astype(np.float32)


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


X43_INT = np.array([[1, 2, 3], [4, 5, 6], [7, 8, 9], [10, 11, 12]], dtype=np.float32)
X43_FRAC = np.array([[1.0, 0.5, 0.25], [0.25, 0.5, 1.0], [1.5, 2.0, 0.0], [0.0, 0.25, 2.0]], dtype=np.float32)

_COMP = {"tall_skinny": (256, 64, 16, False), "short_wide": (64, 256, 16, False),
         "square": (192, 192, 24, False), "rank_deficient": (200, 80, 16, True)}


def cases() -> dict:
    """name -> (kind, X, k, algorithm).  Inputs of the reference's own tests first."""
    c = {}
    c["pca_4x3_int"] = ("pca", X43_INT, 2, "cusolver")
    c["tsvd_4x3_frac"] = ("tsvd", X43_FRAC, 2, "cusolver")
    c["pca_128x32_k8"] = ("pca", np.random.default_rng(0).normal(size=(128, 32)).astype(np.float32), 8, "cusolver")
    c["tsvd_160x40_k10"] = ("tsvd", np.random.default_rng(1).normal(size=(160, 40)).astype(np.float32), 10, "cusolver")
    for name, (n, m, k, rd) in _COMP.items():
        for seed in (0, 7):
            X = comprehensive_matrix(n, m, k, rd, np.float32, seed)
            c[f"pca_{name}_s{seed}"] = ("pca", X, k, "cusolver")
            c[f"tsvd_{name}_s{seed}"] = ("tsvd", X, k, "power")  # min(n,m) <= 256 -> reference takes its exact path
    c["pca_offset_2000x96_k12"] = ("pca", offset_matrix(), 12, "cusolver")
    c["tsvd_spiked_3000x320_k16_exact"] = ("tsvd", spiked_matrix(), 16, "cusolver")
    c["pca_spiked_3000x320_k16_exact"] = ("pca", spiked_matrix(), 16, "cusolver")
    c["tsvd_spiked_3000x320_k16_power"] = ("tsvd", spiked_matrix(), 16, "power")  # reference's randomized path
    c["tsvd_config1_10000x100_k10"] = ("tsvd", config1_matrix(), 10, "power")
    return c
