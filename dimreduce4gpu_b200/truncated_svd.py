"""TruncatedSVD estimator - same surface as the reference (dimreduce4gpu/truncated_svd.py:19-118).

Constructor arguments, ``fit`` / ``fit_transform`` / ``transform`` and the attributes
``components_``, ``singular_values_``, ``explained_variance_``, ``explained_variance_ratio_`` and
the private ``_U`` are preserved.  Additions: ``inverse_transform`` (absent from the reference,
SURVEY fact 8; sklearn semantics), ``transform`` runs on the GPU (reference: host numpy,
truncated_svd.py:116-118), torch CUDA tensors are accepted and stay on the device, and
``dtype="float64"`` keeps doubles (the reference always down-casts, truncated_svd.py:72)."""
from __future__ import annotations

import ctypes
from typing import Literal, Optional

import numpy as np

from ._backend import select_backend
from .lib_dimreduce4gpu import _load_tsvd_lib, params

Backend = Literal["auto", "gpu", "cpu"]

_contexts: dict[int, "object"] = {}


def _device_context(gpu_id: int):
    """One persistent DeviceContext per device for the estimator-level device paths."""
    from .device import DeviceContext

    ctx = _contexts.get(gpu_id)
    if ctx is None:
        ctx = _contexts[gpu_id] = DeviceContext(gpu_id)
    return ctx


def _is_torch_cuda(X) -> bool:
    return type(X).__module__.startswith("torch") and getattr(X, "is_cuda", False)


def _ptr(x: np.ndarray):
    ct = ctypes.c_float if x.dtype == np.float32 else ctypes.c_double
    return x.ctypes.data_as(ctypes.POINTER(ct))


def _is_sparse(X) -> bool:
    return type(X).__module__.startswith("scipy.sparse")


def _densify(X):
    if _is_sparse(X):
        return X.toarray()  # reference densifies CSR on the host (truncated_svd.py:67-70); only tiny inputs come here
    return X


class _DeviceCSR:
    """A scipy sparse matrix as device-resident CSR arrays (int64 indptr, int32 indices, values in `dtype`): the only
    form in which sparse input crosses PCIe.  Row chunks are expanded on the device (csrc/csr.cu) into the dense tiles
    the streamed kernels read; the reference densifies the whole matrix on the host first (truncated_svd.py:67-70)."""

    CHUNK_BYTES = 256 << 20

    def __init__(self, X, dtype, device: int) -> None:
        import torch

        X = X.tocsr()
        X.sum_duplicates()
        self.n, self.d = X.shape
        self.tdt = torch.float32 if dtype == np.float32 else torch.float64
        dev = f"cuda:{device}"
        self.indptr = torch.from_numpy(np.ascontiguousarray(X.indptr, dtype=np.int64)).to(dev)
        self.indices = torch.from_numpy(np.ascontiguousarray(X.indices, dtype=np.int32)).to(dev)
        self.data = torch.from_numpy(np.ascontiguousarray(X.data, dtype=dtype)).to(dev)
        self.rows = max(1, min(self.n, self.CHUNK_BYTES // max(1, self.d * self.data.element_size())))
        self.buf = torch.empty(self.rows, self.d, dtype=self.tdt, device=dev)

    def chunks(self, ctx):
        """(row0, dense tile) per chunk; the tile aliases one buffer, ordered by the stream."""
        for r0 in range(0, self.n, self.rows):
            rows = min(self.rows, self.n - r0)
            yield r0, ctx.csr_to_dense(self.indptr, self.indices, self.data, r0, rows, self.d, out=self.buf[:rows])


class TruncatedSVD:
    """Truncated SVD on the B200 native backend."""

    _center = False
    _entry = "tsvd"

    def __init__(self, n_components: int = 2, algorithm: str = "power", n_iter: int = 5,
                 random_state: Optional[int] = None, tol: float = 1e-5, verbose: bool = False, gpu_id: int = 0,
                 backend: Backend = "auto", *, dtype: Optional[str] = None) -> None:
        self.n_components = int(n_components)
        self.algorithm = str(algorithm)
        self.n_iter = int(n_iter)
        self.random_state = int(random_state) if random_state is not None else int(np.random.randint(0, 2**31 - 1))
        self.tol = float(tol)
        self.verbose = bool(verbose)
        self.gpu_id = int(gpu_id)
        self.backend: Backend = backend
        self.dtype = np.dtype(dtype if dtype is not None else np.float32)
        if self.dtype not in (np.dtype(np.float32), np.dtype(np.float64)):
            raise ValueError("dtype must be float32 or float64")
        self._Q = None
        self._w = None
        self._U = None
        self.explained_variance_ = None
        self.explained_variance_ratio_ = None
        self.mean_ = None

    # ------------------------------------------------------------------ attributes
    @property
    def components_(self):
        if getattr(self, "_stream_dirty", False):
            self.finalize_fit()
        if self._Q is None:
            raise AttributeError("components_ is not available before fit/fit_transform.")
        return self._Q

    @property
    def singular_values_(self):
        if getattr(self, "_stream_dirty", False):
            self.finalize_fit()
        if self._w is None:
            raise AttributeError("singular_values_ is not available before fit/fit_transform.")
        return self._w

    # ------------------------------------------------------------------ fit
    def _params(self, n: int, m: int, k: int) -> params:
        p = params()
        p.X_n, p.X_m, p.k = n, m, k
        p.algorithm = self.algorithm.encode("utf-8")
        p.n_iter, p.random_state, p.tol = self.n_iter, self.random_state, float(self.tol)
        p.verbose, p.gpu_id = (1 if self.verbose else 0), self.gpu_id
        p.whiten = bool(getattr(self, "whiten", False))
        return p

    def fit(self, X, y=None):
        if _is_torch_cuda(X):
            select_backend(self.backend)
            self._drop_stream_state()
            self._fit_device(X, want_scores=False)
        else:
            self.fit_transform(X)
        return self

    def _drop_stream_state(self):
        """A one-shot fit replaces whatever partial_fit accumulated: without this a later read of components_ would
        finalize the stale chunks over the fresh fit."""
        if getattr(self, "_stream_dirty", False) or getattr(self, "_stream_d", None) is not None:
            self.reset_partial_fit()

    def fit_transform(self, X):
        select_backend(self.backend)
        self._drop_stream_state()
        if _is_torch_cuda(X):
            return self._fit_device(X, want_scores=True)
        if _is_sparse(X):
            return self._fit_csr(X)
        X = np.ascontiguousarray(_densify(X), dtype=self.dtype)
        if X.ndim != 2:
            raise ValueError("X must be 2-D (n_samples, n_features)")
        n, m = X.shape
        k = min(self.n_components, n, m)
        dt = self.dtype
        Q = np.zeros((k, m), dtype=dt)
        w = np.zeros((k,), dtype=dt)
        U = np.zeros((n, k), dtype=dt)
        Xt = np.zeros((n, k), dtype=dt)
        ev = np.zeros((k,), dtype=dt)
        evr = np.zeros((k,), dtype=dt)
        mean = np.zeros((m,), dtype=dt)
        p = self._params(n, m, k)
        sfx = "float" if dt == np.float32 else "double"
        from ._native import last_error

        if self._center:
            from .lib_dimreduce4gpu import _load_pca_lib

            _load_pca_lib(sfx)(_ptr(X), _ptr(Q), _ptr(w), _ptr(U), _ptr(Xt), _ptr(ev), _ptr(evr), _ptr(mean), p)
        else:
            _load_tsvd_lib(sfx)(_ptr(X), _ptr(Q), _ptr(w), _ptr(U), _ptr(Xt), _ptr(ev), _ptr(evr), p)
        err = last_error()
        if err:  # the C symbols are void like the reference's; the message is out of band
            raise RuntimeError(f"libdimreduce4gpu: {err}")
        self._Q, self._w, self._U = Q, w, U
        self.explained_variance_, self.explained_variance_ratio_ = ev, evr
        self.mean_ = mean if self._center else None
        return Xt

    def _fit_device(self, X, want_scores: bool):
        """X is a torch CUDA tensor: nothing leaves the device except the k- and d-sized results."""
        ctx = _device_context(X.device.index)
        res = ctx.fit(X, self.n_components, center=self._center, algorithm=self.algorithm, n_iter=self.n_iter,
                      random_state=self.random_state, tol=self.tol, verbose=self.verbose,
                      whiten=bool(getattr(self, "whiten", False)))
        self._Q_dev, self._w_dev, self._mean_dev = res["components"], res["singular_values"], res["mean"]
        Z = None
        if want_scores:
            Z = ctx.transform(X, res["components"], mean=res["mean"])
        self._Q = res["components"].cpu().numpy()
        self._w = res["singular_values"].cpu().numpy()
        self.explained_variance_ = res["explained_variance"].cpu().numpy()
        self.explained_variance_ratio_ = res["explained_variance_ratio"].cpu().numpy()
        self.mean_ = res["mean"].cpu().numpy() if self._center else None
        self._U = None
        return Z

    def _fit_csr(self, X):
        """scipy sparse input: CSR arrays to the device, row chunks expanded there and streamed through
        the fit's accumulators (dr4g_partial_fit_*), then a second sweep for the scores.  Nothing dense exists on the
        host, and on the device only one chunk at a time."""
        import torch

        from .device import DeviceContext

        n, d = X.shape
        k = min(self.n_components, n, d)
        with torch.cuda.device(self.gpu_id):
            ctx = DeviceContext(self.gpu_id)
            csr = _DeviceCSR(X, self.dtype, self.gpu_id)
            for _r0, tile in csr.chunks(ctx):
                ctx.partial_fit(tile, center=self._center)
            res = ctx.partial_fit_finalize(d, k, self._center, csr.tdt, algorithm=self.algorithm, n_iter=self.n_iter,
                                           random_state=self.random_state, tol=self.tol, verbose=self.verbose,
                                           whiten=bool(getattr(self, "whiten", False)))
            Xt = np.empty((n, k), dtype=self.dtype)
            U = np.empty((n, k), dtype=self.dtype)
            for r0, tile in csr.chunks(ctx):
                Z, Uc = ctx.transform(tile, res["components"], mean=res["mean"], singular_values=res["singular_values"],
                                      want_u=True)
                Xt[r0:r0 + tile.shape[0]] = Z.cpu().numpy()
                U[r0:r0 + tile.shape[0]] = Uc.cpu().numpy()
            ctx.sync()
        self._Q = res["components"].cpu().numpy()
        self._w = res["singular_values"].cpu().numpy()
        self.explained_variance_ = res["explained_variance"].cpu().numpy()
        self.explained_variance_ratio_ = res["explained_variance_ratio"].cpu().numpy()
        self.mean_ = res["mean"].cpu().numpy() if self._center else None
        self._U = U
        return Xt

    # ------------------------------------------------------------------ streaming fit
    def partial_fit(self, X, y=None):
        """Accumulate a chunk of rows (numpy array or torch CUDA tensor, (n_chunk, n_features)) into the
        fit's sufficient statistics on the GPU.  X as a whole never has to fit in device memory; call
        ``finalize_fit()`` (or read ``components_`` through ``transform``) after the last chunk.  Not in the
        reference, whose entry points copy all of X to the device (src/tsvd.cu:427-452)."""
        import torch

        select_backend(self.backend)
        tdt = torch.float32 if self.dtype == np.float32 else torch.float64
        Xd, _ = self._to_device(X, tdt)
        if Xd.dim() != 2:
            raise ValueError("X must be 2-D (n_samples, n_features)")
        if getattr(self, "_sctx", None) is None:
            from .device import DeviceContext

            self._sctx = DeviceContext(Xd.device.index)   # the accumulators live in this private context
            self._stream_d = None
        if self._stream_d is None:
            self._sctx.partial_fit_reset()
            self._stream_d = int(Xd.shape[1])
        elif Xd.shape[1] != self._stream_d:
            raise ValueError(f"chunk has {Xd.shape[1]} features, earlier chunks had {self._stream_d}")
        with torch.cuda.device(Xd.device):
            self._sctx.partial_fit(Xd, center=self._center)
        self._stream_dirty = True
        return self

    def finalize_fit(self):
        """Eigensolve + outputs from the chunks accumulated by partial_fit (more chunks may follow)."""
        import torch

        if getattr(self, "_sctx", None) is None or self._stream_d is None:
            raise RuntimeError("finalize_fit() needs at least one partial_fit() call")
        tdt = torch.float32 if self.dtype == np.float32 else torch.float64
        with torch.cuda.device(self._sctx.device):
            res = self._sctx.partial_fit_finalize(self._stream_d, self.n_components, self._center, tdt,
                                                  algorithm=self.algorithm, n_iter=self.n_iter,
                                                  random_state=self.random_state, tol=self.tol, verbose=self.verbose,
                                                  whiten=bool(getattr(self, "whiten", False)))
            self._sctx.sync()
        self._Q = res["components"].cpu().numpy()
        self._w = res["singular_values"].cpu().numpy()
        self.explained_variance_ = res["explained_variance"].cpu().numpy()
        self.explained_variance_ratio_ = res["explained_variance_ratio"].cpu().numpy()
        self.mean_ = res["mean"].cpu().numpy() if self._center else None
        self._U = None
        self._stream_dirty = False
        return self

    def reset_partial_fit(self):
        if getattr(self, "_sctx", None) is not None:
            self._sctx.partial_fit_reset()
            self._stream_d = None
        self._stream_dirty = False
        return self

    # ------------------------------------------------------------------ transform / inverse
    def _to_device(self, A, dtype):
        import torch

        if _is_torch_cuda(A):
            return A.contiguous().to(dtype), True
        a = np.ascontiguousarray(_densify(A), dtype=np.float32 if dtype == torch.float32 else np.float64)
        return torch.from_numpy(a).to(f"cuda:{self.gpu_id}"), False

    def transform(self, X, center: bool = False):
        """X @ components_.T on the GPU.  ``center=False`` reproduces the reference, whose transform
        never subtracts the mean, not even for PCA (truncated_svd.py:116-118); ``center=True``
        gives the sklearn behaviour (X - mean_) @ components_.T."""
        import torch

        select_backend(self.backend)
        Q = self.components_
        tdt = torch.float32 if Q.dtype == np.float32 else torch.float64
        if _is_sparse(X):
            ctx = _device_context(self.gpu_id)
            with torch.cuda.device(self.gpu_id):
                csr = _DeviceCSR(X, Q.dtype, self.gpu_id)
                Qd = torch.from_numpy(np.ascontiguousarray(Q)).to(csr.data.device)
                md = torch.from_numpy(np.ascontiguousarray(self.mean_)).to(csr.data.device) if center and self.mean_ is not None else None
                out = np.empty((csr.n, Q.shape[0]), dtype=Q.dtype)
                for r0, tile in csr.chunks(ctx):
                    out[r0:r0 + tile.shape[0]] = ctx.transform(tile, Qd, mean=md).cpu().numpy()
            return out
        Xd, on_dev = self._to_device(X, tdt)
        ctx = _device_context(Xd.device.index)
        Qd = torch.from_numpy(np.ascontiguousarray(Q)).to(Xd.device)
        md = None
        if center and self.mean_ is not None:
            md = torch.from_numpy(np.ascontiguousarray(self.mean_)).to(Xd.device)
        Z = ctx.transform(Xd, Qd, mean=md)
        return Z if on_dev else Z.cpu().numpy()

    def inverse_transform(self, Z):
        """Z @ components_ (+ mean_ for PCA).  Not in the reference; sklearn semantics."""
        import torch

        select_backend(self.backend)
        Q = self.components_
        tdt = torch.float32 if Q.dtype == np.float32 else torch.float64
        Zd, on_dev = self._to_device(Z, tdt)
        if getattr(self, "whiten", False):
            # components_ rows carry 1/sqrt(ev): Xc = sum_i z_i ev_i q_i
            Zd = Zd * torch.from_numpy(np.ascontiguousarray(self.explained_variance_)).to(Zd.device, Zd.dtype)
        ctx = _device_context(Zd.device.index)
        Qd = torch.from_numpy(np.ascontiguousarray(Q)).to(Zd.device)
        md = torch.from_numpy(np.ascontiguousarray(self.mean_)).to(Zd.device) if self.mean_ is not None else None
        X = ctx.inverse_transform(Zd, Qd, mean=md)
        return X if on_dev else X.cpu().numpy()
