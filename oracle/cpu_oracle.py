"""CPU oracle for the PCA / TruncatedSVD fit+transform hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import this module.  The product (``dimreduce4gpu_b200``) never does;
it has no CPU fallback.

Two oracles live here:

1. ``RefLib`` - a ctypes binding of ``oracle/_ref/libdimreduce4cpu.so``, i.e. the reference's
   own CPU backend compiled unmodified from ``/root/reference/src/cpu_backend.cpp`` by
   ``oracle/Makefile``.  This *is* the reference; parity claims are anchored on it.
2. ``restated_*`` - a numpy/LAPACK restatement of the same algorithm, each function citing
   the reference lines it follows.  It exists (a) to run in float64, which the reference CPU
   backend cannot (``include/cpu_backend.h:32-51`` has float entry points only and the wrapper
   down-casts, ``dimreduce4gpu/truncated_svd.py:72``), and (b) as a portable checker when the
   ``_ref`` library is not present.

Parity pinning: the float32 restatement is checked against ``RefLib`` on every golden input
in ``tests/golden/`` (``tests/test_oracle.py``); the goldens themselves are outputs of the
reference run in the build container (``tests/golden/make_golden.py``).  The float64
restatement has no reference implementation to pin against ("parity unpinned" at fp64 by the
reference itself) beyond agreeing with the float32 reference to float32 accuracy.
"""
from __future__ import annotations

import ctypes
import os
from dataclasses import dataclass
from pathlib import Path
from typing import Optional

import numpy as np

_HERE = Path(__file__).resolve().parent
REF_SO = _HERE / "_ref" / "libdimreduce4cpu.so"


# --------------------------------------------------------------------------------------
# params struct: include/cpu_backend.h:19-30 (same layout as include/tsvd.h:24-35)
# --------------------------------------------------------------------------------------
class _Params(ctypes.Structure):
    _fields_ = [
        ("X_n", ctypes.c_int32),
        ("X_m", ctypes.c_int32),
        ("k", ctypes.c_int32),
        ("algorithm", ctypes.c_char_p),
        ("n_iter", ctypes.c_int32),
        ("random_state", ctypes.c_int32),
        ("tol", ctypes.c_float),
        ("verbose", ctypes.c_int32),
        ("gpu_id", ctypes.c_int32),
        ("whiten", ctypes.c_bool),
    ]


@dataclass
class FitResult:
    components: np.ndarray  # (k, d)   == Q
    singular_values: np.ndarray  # (k,) == w
    U: np.ndarray  # (n, k)
    X_transformed: np.ndarray  # (n, k)
    explained_variance: np.ndarray  # (k,)
    explained_variance_ratio: np.ndarray  # (k,)
    mean: Optional[np.ndarray] = None  # (d,) PCA only


def _fp(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


class RefLib:
    """The reference CPU backend itself (src/cpu_backend.cpp:336-389) through ctypes."""

    def __init__(self, path: os.PathLike | str | None = None) -> None:
        p = Path(path) if path is not None else REF_SO
        if not p.is_file():
            raise FileNotFoundError(f"{p} missing - run `make -C oracle` where /root/reference exists")
        self.path = p
        self._lib = ctypes.CDLL(str(p))
        f7 = [ctypes.POINTER(ctypes.c_float)] * 7
        self._lib.truncated_svd_float.argtypes = f7 + [_Params]
        self._lib.truncated_svd_float.restype = None
        self._lib.pca_float.argtypes = f7 + [ctypes.POINTER(ctypes.c_float), _Params]
        self._lib.pca_float.restype = None

    @staticmethod
    def set_threads(n: int) -> None:
        """OpenBLAS thread count of the scipy-bundled library the oracle links."""
        import glob

        import scipy

        sp = Path(scipy.__file__).resolve().parent.parent / "scipy.libs"
        for so in glob.glob(str(sp / "libscipy_openblas-*.so")):
            lib = ctypes.CDLL(so)
            if hasattr(lib, "scipy_openblas_set_num_threads"):
                lib.scipy_openblas_set_num_threads(ctypes.c_int(int(n)))

    def _run(self, X: np.ndarray, k: int, algorithm: str, n_iter: int, random_state: int, pca: bool) -> FitResult:
        X = np.ascontiguousarray(X, dtype=np.float32)
        n, d = X.shape
        k = min(int(k), n, d)
        Q = np.zeros((k, d), np.float32)
        w = np.zeros((k,), np.float32)
        U = np.zeros((n, k), np.float32)
        Xt = np.zeros((n, k), np.float32)
        ev = np.zeros((k,), np.float32)
        evr = np.zeros((k,), np.float32)
        mean = np.zeros((d,), np.float32)
        alg = algorithm.encode()
        p = _Params(n, d, k, alg, int(n_iter), int(random_state), 1e-5, 0, 0, False)
        if pca:
            self._lib.pca_float(_fp(X), _fp(Q), _fp(w), _fp(U), _fp(Xt), _fp(ev), _fp(evr), _fp(mean), p)
        else:
            self._lib.truncated_svd_float(_fp(X), _fp(Q), _fp(w), _fp(U), _fp(Xt), _fp(ev), _fp(evr), p)
        return FitResult(Q, w, U, Xt, ev, evr, mean if pca else None)

    def pca(self, X, k, algorithm="cusolver", n_iter=5, random_state=0) -> FitResult:
        return self._run(X, k, algorithm, n_iter, random_state, True)

    def tsvd(self, X, k, algorithm="power", n_iter=5, random_state=0) -> FitResult:
        return self._run(X, k, algorithm, n_iter, random_state, False)


# --------------------------------------------------------------------------------------
# numpy restatement
# --------------------------------------------------------------------------------------
def _gesdd_topk(A: np.ndarray, k: int):
    """exact_svd_topk_colmajor, src/cpu_backend.cpp:72-135: LAPACK ?gesdd(jobz='S') on a copy
    of the (centred) matrix, keep the leading k triplets (gesvd on failure, :97-110)."""
    import scipy.linalg

    try:
        U, s, Vt = scipy.linalg.svd(A, full_matrices=False, lapack_driver="gesdd", check_finite=False)
    except np.linalg.LinAlgError:
        U, s, Vt = scipy.linalg.svd(A, full_matrices=False, lapack_driver="gesvd", check_finite=False)
    return U[:, :k], s[:k], Vt[:k]


def _explained_variance(X_for_var: np.ndarray, s: np.ndarray):
    """compute_explained_variance_rowmajor, src/cpu_backend.cpp:277-300.
    ev = s^2 / max(1, n-1)               (:279-282, double arithmetic, cast to T)
    total_var = sum_j var_j(ddof -> max(1, n-1)) about the column mean, in double (:283-295)
    total_var <= 0 -> 1                  (:296)
    evr = ev (already rounded to T) / total_var   (:297-299)"""
    n = X_for_var.shape[0]
    denom = float(max(1, n - 1))
    dt = X_for_var.dtype
    ev = ((s.astype(np.float64) ** 2) / denom).astype(dt)
    Xd = X_for_var.astype(np.float64)
    mu = Xd.sum(axis=0) / float(n)
    total = float((((Xd - mu) ** 2).sum(axis=0) / denom).sum())
    if total <= 0.0:
        total = 1.0
    evr = (ev.astype(np.float64) / total).astype(dt)
    return ev, evr


def _outputs(U, s, Vt, dtype) -> tuple:
    """fill_outputs_rowmajor, src/cpu_backend.cpp:302-332: w = S; Q = VT (k x d row-major, no
    svd_flip); U (n x k); X_transformed = U * w (product formed in T, :329)."""
    U = np.ascontiguousarray(U, dtype=dtype)
    s = np.ascontiguousarray(s, dtype=dtype)
    Q = np.ascontiguousarray(Vt, dtype=dtype)
    Xt = (U * s[None, :]).astype(dtype)
    return Q, s, U, Xt


def restated_tsvd(X: np.ndarray, k: int, dtype=np.float32) -> FitResult:
    """truncated_svd_float with the exact path, src/cpu_backend.cpp:338-357.  (The exact path
    is taken for algorithm=="cusolver" or min(n,d) <= 256, :347.)  Variance is about the
    column mean of the *uncentred* X (:355 passes X itself)."""
    X = np.ascontiguousarray(X, dtype=dtype)
    n, d = X.shape
    k = min(int(k), n, d)  # :342
    U, s, Vt = _gesdd_topk(X, k)
    Q, w, U, Xt = _outputs(U, s, Vt, dtype)
    ev, evr = _explained_variance(X, w)
    return FitResult(Q, w, U, Xt, ev, evr, None)


def restated_pca(X: np.ndarray, k: int, dtype=np.float32) -> FitResult:
    """pca_float with the exact path, src/cpu_backend.cpp:359-387.
    mean accumulated in double, stored as T (:54-60); centring X - mean done in double against
    the double mean then rounded to T (:63-68); variance is taken on X - mean_T formed in T
    (:379-385)."""
    X = np.ascontiguousarray(X, dtype=dtype)
    n, d = X.shape
    k = min(int(k), n, d)  # :363
    mean_d = X.astype(np.float64).sum(axis=0) / float(n)
    mean = mean_d.astype(dtype)
    Xc = (X.astype(np.float64) - mean_d).astype(dtype)
    U, s, Vt = _gesdd_topk(Xc, k)
    Q, w, U, Xt = _outputs(U, s, Vt, dtype)
    Xc_row = (X - mean[None, :]).astype(dtype)
    ev, evr = _explained_variance(Xc_row, w)
    return FitResult(Q, w, U, Xt, ev, evr, mean)


def restated_randomized_tsvd(X: np.ndarray, k: int, n_iter: int, random_state: int, dtype=np.float32,
                             center: bool = False) -> FitResult:
    """randomized_svd_topk_colmajor, src/cpu_backend.cpp:167-275 (taken when algorithm !=
    "cusolver" and min(n,d) > 256, :347/:369).  Omega ~ N(0,1) d x l with l = min(k+10, min(n,d)),
    Y = X Omega; n_iter x { Z = X^T Y; Y = X Z; Y = qr(Y).Q }; Q = qr(Y).Q; B = Q^T X;
    gesdd(B); U = Q Uhat[:, :k].  The reference draws Omega from std::mt19937 +
    std::normal_distribution (:173-178), a stream numpy cannot reproduce, so this restatement
    matches the reference only statistically (same algorithm, different Omega)."""
    X = np.ascontiguousarray(X, dtype=dtype)
    n, d = X.shape
    k = min(int(k), n, d)
    mean = None
    Xw = X
    if center:
        mean_d = X.astype(np.float64).sum(axis=0) / float(n)
        mean = mean_d.astype(dtype)
        Xw = (X.astype(np.float64) - mean_d).astype(dtype)
    l = min(k + 10, min(n, d))
    rng = np.random.default_rng(12345 if random_state <= 0 else random_state)
    Om = rng.standard_normal((d, l)).astype(dtype)
    Y = Xw @ Om
    for _ in range(max(0, n_iter)):
        Z = Xw.T @ Y
        Y = Xw @ Z
        Y, _ = np.linalg.qr(Y)
    Qb, _ = np.linalg.qr(Y)
    B = Qb.T @ Xw
    Uh, s, Vt = _gesdd_topk(B, k)
    U = Qb @ Uh
    Q, w, U, Xt = _outputs(U, s, Vt, dtype)
    Xvar = (X - mean[None, :]).astype(dtype) if center else X
    ev, evr = _explained_variance(Xvar, w)
    return FitResult(Q, w, U, Xt, ev, evr, mean)


# --------------------------------------------------------------------------------------
# comparison helpers shared by the parity tests
# --------------------------------------------------------------------------------------
def svd_flip_v(components: np.ndarray, *others: np.ndarray):
    """Deterministic v-based sign convention applied to BOTH sides before comparing
    (the reference CPU backend applies no flip, src/cpu_backend.cpp:302-332, so its signs are
    whatever LAPACK returns): make the largest-|.| entry of each component row positive.
    `others` are (n, k) arrays whose columns are flipped with the same signs."""
    C = np.array(components, copy=True)
    idx = np.argmax(np.abs(C), axis=1)
    sg = np.sign(C[np.arange(C.shape[0]), idx])
    sg[sg == 0] = 1
    C *= sg[:, None]
    outs = [np.array(o, copy=True) * sg[None, :] for o in others]
    return (C, *outs)


def max_principal_angle(C1: np.ndarray, C2: np.ndarray) -> float:
    """Largest principal angle (radians) between the row spaces of C1 and C2."""
    Q1, _ = np.linalg.qr(np.asarray(C1, np.float64).T)
    Q2, _ = np.linalg.qr(np.asarray(C2, np.float64).T)
    # sin-based formula is accurate for small angles
    R = Q2 - Q1 @ (Q1.T @ Q2)
    s = np.linalg.svd(R, compute_uv=False)
    return float(np.arcsin(np.clip(s.max(), 0.0, 1.0)))
