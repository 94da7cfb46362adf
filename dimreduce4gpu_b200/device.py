"""Device-resident, row-sharded operation on top of the dr4g_* symbols of include/dimreduce4gpu.h.

The reference has no such layer: it uploads X on every call (src/tsvd.cu:430-431) and is single
device (src/tsvd.cu:429).  Here X stays in HBM as a torch CUDA tensor (torch is only the
allocator / stream / rendezvous plumbing; every kernel is in libdimreduce4gpu.so).

One process per GPU: ``DeviceContext.init_comm_from_torch()`` ships rank 0's NCCL unique id through
``torch.distributed`` and gives the context its own communicator; ``fit`` then all-reduces the
packed [Gram | column sums | sums of squares | row count] once."""
from __future__ import annotations

import ctypes
import functools
import threading
from typing import Optional

import numpy as np

from ._native import last_error, load_library
from .lib_dimreduce4gpu import params


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed: {last_error()}")


def _suffix(t) -> str:
    import torch

    if t.dtype == torch.float32:
        return "float"
    if t.dtype == torch.float64:
        return "double"
    raise TypeError(f"float32 or float64 expected, got {t.dtype}")


def _need_cuda_contig(t, name: str, device: int):
    if not t.is_cuda or t.device.index != device:
        raise ValueError(f"{name} must live on cuda:{device}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be C-contiguous (row-major)")
    return t


def _serialized(fn):
    """Stream selection and launch are one critical section per context: the estimators share one cached context per
    GPU, and dr4g_ctx_set_stream followed by a launch from two Python threads could otherwise interleave (the C entry
    points take the context's mutex one by one)."""
    @functools.wraps(fn)
    def wrapper(self, *a, **kw):
        with self._lock:
            return fn(self, *a, **kw)
    return wrapper


class DeviceContext:
    """Persistent per-device context (stream, workspace arena, optional NCCL communicator)."""

    def __init__(self, device: int = 0) -> None:
        self._lib = load_library()
        self.device = int(device)
        self._h = self._lib.dr4g_ctx_create(self.device)
        if not self._h:
            raise RuntimeError(f"dr4g_ctx_create({device}) failed: {last_error()}")
        self.rank, self.world = 0, 1
        self._lock = threading.RLock()

    # ------------------------------------------------------------------ lifecycle
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.dr4g_ctx_destroy(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _on_torch_stream(self) -> None:
        import torch

        s = torch.cuda.current_stream(self.device).cuda_stream
        if s == 0:
            s = 1  # torch's default stream is the legacy NULL stream; NULL means "private stream" in the C ABI -> cudaStreamLegacy
        _check(self._lib.dr4g_ctx_set_stream(self._h, ctypes.c_void_p(s)), "dr4g_ctx_set_stream")

    def set_option(self, key: str, value: float) -> None:
        _check(self._lib.dr4g_set_option(self._h, key.encode(), float(value)), f"dr4g_set_option({key})")

    def sync(self) -> None:
        _check(self._lib.dr4g_sync(self._h), "dr4g_sync")

    # ------------------------------------------------------------------ multi-GPU
    def init_comm(self, unique_id: bytes, rank: int, world: int) -> None:
        buf = ctypes.create_string_buffer(bytes(unique_id), 128)
        _check(self._lib.dr4g_comm_init(self._h, buf, int(rank), int(world)), "dr4g_comm_init")
        self.rank, self.world = int(rank), int(world)

    def new_unique_id(self) -> bytes:
        buf = ctypes.create_string_buffer(128)
        _check(self._lib.dr4g_nccl_unique_id(buf), "dr4g_nccl_unique_id")
        return buf.raw

    def init_comm_from_torch(self) -> None:
        """Rendezvous through an initialised torch.distributed group (any backend)."""
        import torch
        import torch.distributed as dist

        rank, world = dist.get_rank(), dist.get_world_size()
        if world == 1:
            return
        uid = [self.new_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        with torch.cuda.device(self.device):
            self.init_comm(uid[0], rank, world)
            self._init_p2p(dist, rank, world)

    def _init_p2p(self, dist, rank: int, world: int) -> None:
        """Peer-memory exchange (csrc/p2p.cu) when every rank can map every other rank's buffer (one NVSwitch box);
        any failure on any rank leaves all ranks on NCCL.  DR4G_P2P_MAX_D (default 1024, 0 = off) sizes the buffers."""
        import os

        max_d = int(os.environ.get("DR4G_P2P_MAX_D", "1024"))
        handle = ctypes.create_string_buffer(64)
        ok = 0 < max_d and 2 <= world <= 16
        if ok:
            ok = self._lib.dr4g_p2p_prepare(self._h, max_d * max_d + 3 * max_d + 1, handle) == 0
        got = [None] * world
        dist.all_gather_object(got, (bool(ok), handle.raw))
        if not all(g[0] for g in got):
            return
        ok = self._lib.dr4g_p2p_connect(self._h, b"".join(g[1] for g in got)) == 0
        got = [None] * world
        dist.all_gather_object(got, bool(ok))
        if not all(got):
            self.set_option("exchange", 1)   # somebody could not map a peer: everybody stays on NCCL

    # ------------------------------------------------------------------ fit / transform
    @_serialized
    def fit(self, X, n_components: int, center: bool, algorithm: str = "cusolver", n_iter: int = 5,
            random_state: int = 0, tol: float = 1e-5, verbose: bool = False, out: Optional[dict] = None,
            whiten: bool = False) -> dict:
        """X: (n_local, d) CUDA tensor, this rank's row shard.  Returns device tensors
        components (k,d), singular_values, explained_variance, explained_variance_ratio (k,), mean (d,)|None.
        Asynchronous on the current torch stream."""
        import torch

        _need_cuda_contig(X, "X", self.device)
        if X.dim() != 2:
            raise ValueError("X must be 2-D (n_samples, n_features)")
        n, d = X.shape
        k = min(int(n_components), d) if self.world > 1 else min(int(n_components), n, d)
        sfx = _suffix(X)
        if out is None:
            z = lambda *s: torch.empty(*s, dtype=X.dtype, device=X.device)  # noqa: E731
            out = dict(components=z(k, d), singular_values=z(k), explained_variance=z(k),
                       explained_variance_ratio=z(k), mean=z(d) if center else None)
        p = params()
        p.X_n, p.X_m, p.k = min(n, 2**31 - 1), d, k
        p.algorithm = algorithm.encode("utf-8")
        p.n_iter, p.random_state, p.tol = int(n_iter), int(random_state) & 0x7FFFFFFF, float(tol)
        p.verbose, p.gpu_id, p.whiten = int(bool(verbose)), self.device, bool(whiten)
        self._on_torch_stream()
        mean_ptr = out["mean"].data_ptr() if center else None
        rc = getattr(self._lib, f"dr4g_fit_{sfx}")(
            self._h, X.data_ptr(), n, d, k, 1 if center else 0, p, out["components"].data_ptr(),
            out["singular_values"].data_ptr(), out["explained_variance"].data_ptr(),
            out["explained_variance_ratio"].data_ptr(), mean_ptr)
        _check(rc, f"dr4g_fit_{sfx}")
        return out

    @_serialized
    def fit_transform_host(self, X: np.ndarray, n_components: int, center: bool, algorithm: str = "cusolver",
                           n_iter: int = 5, random_state: int = 0, tol: float = 1e-5, verbose: bool = False,
                           whiten: bool = False, want_u: bool = True, out: Optional[dict] = None) -> dict:
        """Host-resident row shard (numpy, C-contiguous fp32/fp64; pinned or pageable) -> model + this shard's scores
        in host arrays (dr4g_fit_transform_host_*).  With a communicator the fit is the row-sharded one.  Blocking."""
        if X.ndim != 2 or not X.flags.c_contiguous or X.dtype not in (np.float32, np.float64):
            raise ValueError("X must be a C-contiguous 2-D float32/float64 array")
        n, d = X.shape
        k = min(int(n_components), d) if self.world > 1 else min(int(n_components), n, d)
        sfx = "float" if X.dtype == np.float32 else "double"
        if out is None:
            z = lambda *s: np.zeros(s, dtype=X.dtype)  # noqa: E731
            out = dict(components=z(k, d), singular_values=z(k), explained_variance=z(k), explained_variance_ratio=z(k),
                       mean=z(d) if center else None, X_transformed=z(n, k), U=z(n, k) if want_u else None)
        p = params()
        p.X_n, p.X_m, p.k = min(n, 2**31 - 1), d, k
        p.algorithm = algorithm.encode("utf-8")
        p.n_iter, p.random_state, p.tol = int(n_iter), int(random_state) & 0x7FFFFFFF, float(tol)
        p.verbose, p.gpu_id, p.whiten = int(bool(verbose)), self.device, bool(whiten)
        ptr = lambda a: a.ctypes.data if a is not None else None  # noqa: E731
        rc = getattr(self._lib, f"dr4g_fit_transform_host_{sfx}")(
            self._h, ptr(X), n, d, k, 1 if center else 0, p, ptr(out["components"]), ptr(out["singular_values"]),
            ptr(out.get("U")), ptr(out["X_transformed"]), ptr(out["explained_variance"]),
            ptr(out["explained_variance_ratio"]), ptr(out["mean"]) if center else None)
        _check(rc, f"dr4g_fit_transform_host_{sfx}")
        return out

    # ------------------------------------------------------------------ streaming fit
    @_serialized
    def partial_fit(self, X, center: bool) -> None:
        """Accumulate one device-resident row chunk (n, d) into the context's [G|s|q|n]; asynchronous on the
        current torch stream (keep X alive until later work on that stream has finished)."""
        _need_cuda_contig(X, "X", self.device)
        if X.dim() != 2:
            raise ValueError("X must be 2-D (n_samples, n_features)")
        n, d = X.shape
        sfx = _suffix(X)
        self._on_torch_stream()
        rc = getattr(self._lib, f"dr4g_partial_fit_{sfx}")(self._h, X.data_ptr(), n, d, 1 if center else 0)
        _check(rc, f"dr4g_partial_fit_{sfx}")

    def partial_fit_rows(self) -> int:
        return int(self._lib.dr4g_partial_fit_rows(self._h))

    def partial_fit_reset(self) -> None:
        _check(self._lib.dr4g_partial_fit_reset(self._h), "dr4g_partial_fit_reset")

    @_serialized
    def partial_fit_finalize(self, n_features: int, n_components: int, center: bool, dtype, algorithm: str = "cusolver",
                             n_iter: int = 5, random_state: int = 0, tol: float = 1e-5, verbose: bool = False,
                             whiten: bool = False) -> dict:
        """Eigensolve + outputs from the accumulated chunks; same result dictionary as fit()."""
        import torch

        d = int(n_features)
        rows = self.partial_fit_rows()
        if rows <= 0 and self.world <= 1:
            raise RuntimeError("partial_fit_finalize: nothing accumulated (call partial_fit first)")
        k = min(int(n_components), d) if self.world > 1 else min(int(n_components), rows, d)
        dev = torch.device("cuda", self.device)
        z = lambda *s: torch.empty(*s, dtype=dtype, device=dev)  # noqa: E731
        out = dict(components=z(k, d), singular_values=z(k), explained_variance=z(k),
                   explained_variance_ratio=z(k), mean=z(d) if center else None)
        sfx = "float" if dtype == torch.float32 else "double"
        p = params()
        p.X_n, p.X_m, p.k = min(rows, 2**31 - 1), d, k
        p.algorithm = algorithm.encode("utf-8")
        p.n_iter, p.random_state, p.tol = int(n_iter), int(random_state) & 0x7FFFFFFF, float(tol)
        p.verbose, p.gpu_id, p.whiten = int(bool(verbose)), self.device, bool(whiten)
        self._on_torch_stream()
        rc = getattr(self._lib, f"dr4g_partial_fit_finalize_{sfx}")(
            self._h, k, p, out["components"].data_ptr(), out["singular_values"].data_ptr(),
            out["explained_variance"].data_ptr(), out["explained_variance_ratio"].data_ptr(),
            out["mean"].data_ptr() if center else None)
        _check(rc, f"dr4g_partial_fit_finalize_{sfx}")
        return out

    @_serialized
    def transform(self, X, components, mean=None, singular_values=None, want_u: bool = False, out=None, out_u=None):
        """Z = (X - mean) @ components.T on this rank's rows (no communication); optionally U = Z / sigma."""
        import torch

        _need_cuda_contig(X, "X", self.device)
        _need_cuda_contig(components, "components", self.device)
        n, d = X.shape
        k = components.shape[0]
        if components.shape[1] != d or components.dtype != X.dtype:
            raise ValueError("components must be (k, n_features) with X's dtype")
        sfx = _suffix(X)
        Z = out if out is not None else torch.empty(n, k, dtype=X.dtype, device=X.device)
        U = None
        if want_u:
            if singular_values is None:
                raise ValueError("want_u needs singular_values")
            U = out_u if out_u is not None else torch.empty(n, k, dtype=X.dtype, device=X.device)
        self._on_torch_stream()
        rc = getattr(self._lib, f"dr4g_transform_{sfx}")(
            self._h, X.data_ptr(), n, d, k, components.data_ptr(), mean.data_ptr() if mean is not None else None,
            singular_values.data_ptr() if want_u else None, Z.data_ptr(), U.data_ptr() if want_u else None)
        _check(rc, f"dr4g_transform_{sfx}")
        return (Z, U) if want_u else Z

    @_serialized
    def inverse_transform(self, Z, components, mean=None, out=None):
        import torch

        _need_cuda_contig(Z, "Z", self.device)
        _need_cuda_contig(components, "components", self.device)
        n, k = Z.shape
        d = components.shape[1]
        if components.shape[0] != k or components.dtype != Z.dtype:
            raise ValueError("components must be (k, n_features) with Z's dtype")
        sfx = _suffix(Z)
        X = out if out is not None else torch.empty(n, d, dtype=Z.dtype, device=Z.device)
        self._on_torch_stream()
        rc = getattr(self._lib, f"dr4g_inverse_transform_{sfx}")(
            self._h, Z.data_ptr(), n, d, k, components.data_ptr(), mean.data_ptr() if mean is not None else None,
            X.data_ptr())
        _check(rc, f"dr4g_inverse_transform_{sfx}")
        return X

    @_serialized
    def csr_to_dense(self, indptr, indices, data, row0: int, rows: int, d: int, out=None):
        """Rows [row0, row0 + rows) of a device-resident CSR matrix (int64 indptr, int32 indices, fp32/fp64 data CUDA
        tensors) expanded into a dense (rows, d) CUDA tensor (dr4g_csr_to_dense_*)."""
        import torch

        for t, nm in ((indptr, "indptr"), (indices, "indices"), (data, "data")):
            _need_cuda_contig(t, nm, self.device)
        if indptr.dtype != torch.int64 or indices.dtype != torch.int32:
            raise TypeError("indptr must be int64 and indices int32")
        D = out if out is not None else torch.empty(rows, d, dtype=data.dtype, device=data.device)
        self._on_torch_stream()
        rc = getattr(self._lib, f"dr4g_csr_to_dense_{_suffix(data)}")(
            self._h, indptr.data_ptr(), indices.data_ptr(), data.data_ptr(), int(row0), int(rows), int(d), D.data_ptr())
        _check(rc, "dr4g_csr_to_dense")
        return D

    # ------------------------------------------------------------------ building blocks
    @_serialized
    def colstats(self, X, shift=None):
        import torch

        _need_cuda_contig(X, "X", self.device)
        n, d = X.shape
        s = torch.empty(d, dtype=torch.float64, device=X.device)
        q = torch.empty(d, dtype=torch.float64, device=X.device)
        self._on_torch_stream()
        rc = getattr(self._lib, f"dr4g_colstats_{_suffix(X)}")(
            self._h, X.data_ptr(), n, d, shift.data_ptr() if shift is not None else None, s.data_ptr(), q.data_ptr())
        _check(rc, "dr4g_colstats")
        return s, q

    @_serialized
    def gram(self, X, shift=None):
        import torch

        _need_cuda_contig(X, "X", self.device)
        n, d = X.shape
        G = torch.empty(d, d, dtype=torch.float64, device=X.device)
        self._on_torch_stream()
        rc = getattr(self._lib, f"dr4g_gram_{_suffix(X)}")(
            self._h, X.data_ptr(), n, d, shift.data_ptr() if shift is not None else None, G.data_ptr())
        _check(rc, "dr4g_gram")
        return G

    @_serialized
    def eigh(self, A):
        """Symmetric eigendecomposition (fp64, blocked Jacobi).  Returns (w unsorted, V columns)."""
        import torch

        _need_cuda_contig(A, "A", self.device)
        if A.dtype != torch.float64 or A.dim() != 2 or A.shape[0] != A.shape[1]:
            raise ValueError("A must be a square float64 matrix")
        d = A.shape[0]
        work = A.clone()
        w = torch.empty(d, dtype=torch.float64, device=A.device)
        V = torch.empty(d, d, dtype=torch.float64, device=A.device)
        self._on_torch_stream()
        _check(self._lib.dr4g_eigh_double(self._h, work.data_ptr(), d, w.data_ptr(), V.data_ptr()), "dr4g_eigh_double")
        return w, V

    @_serialized
    def eigh_topk(self, A, k: int):
        """k largest eigenpairs of a symmetric fp64 matrix (the exact fit's solver).  Returns
        (w descending (k,), Vt (k, d) unit eigenvectors in rows)."""
        import torch

        _need_cuda_contig(A, "A", self.device)
        if A.dtype != torch.float64 or A.dim() != 2 or A.shape[0] != A.shape[1]:
            raise ValueError("A must be a square float64 matrix")
        d = A.shape[0]
        work = A.clone()
        w = torch.empty(k, dtype=torch.float64, device=A.device)
        Vt = torch.empty(k, d, dtype=torch.float64, device=A.device)
        self._on_torch_stream()
        _check(self._lib.dr4g_eigh_topk_double(self._h, work.data_ptr(), d, int(k), w.data_ptr(), Vt.data_ptr()),
               "dr4g_eigh_topk_double")
        return w, Vt

    def eigh_topk_static(self, work, k: int, w, Vt) -> None:
        """dr4g_eigh_topk_static_double: eigh_topk with the workspace in the library's static device array - nothing is
        allocated, here or in the library.  `work` (d x d float64, d <= 256) is DESTROYED; `w` (k,) and `Vt` (k, d) are
        caller-provided float64 CUDA tensors.  Calls that overlap on the GPU share the workspace and race."""
        for name, t in (("work", work), ("w", w), ("Vt", Vt)):
            _need_cuda_contig(t, name, self.device)
        d = work.shape[0]
        if work.dim() != 2 or work.shape[1] != d or tuple(w.shape) != (k,) or tuple(Vt.shape) != (k, d):
            raise ValueError("work must be (d, d), w (k,), Vt (k, d)")
        if str(work.dtype) != "torch.float64" or str(w.dtype) != "torch.float64" or str(Vt.dtype) != "torch.float64":
            raise ValueError("work, w, Vt must be float64")
        self._on_torch_stream()
        _check(self._lib.dr4g_eigh_topk_static_double(self._h, work.data_ptr(), d, int(k), w.data_ptr(), Vt.data_ptr()),
               "dr4g_eigh_topk_static_double")

    # ------------------------------------------------------------------ telemetry
    def stage_times(self) -> Optional[dict]:
        """Per-stage device ms of the last fit (call after sync)."""
        buf = (ctypes.c_double * 6)()
        m = self._lib.dr4g_stage_times(self._h, buf, 6)
        if m < 6:
            return None
        names = ("pilot", "gram", "reduce_center", "eigensolve", "outputs", "total")
        return dict(zip(names, np.frombuffer(buf, dtype=np.float64).tolist()))

    @property
    def last_gram_path(self) -> str:
        return {0: "none", 1: "generic_dmma", 2: "tcgen05", 3: "tcgen05_pair", 4: "tcgen05_pair_bf16x3", 5: "dmma_f64_128"}[int(self._lib.dr4g_last_gram_path(self._h))]

    @property
    def last_exchange_path(self) -> str:
        return {0: "none", 1: "nccl", 2: "peer_memory"}[int(self._lib.dr4g_last_exchange_path(self._h))]

    @property
    def last_fold(self) -> int:
        """Row-fold factor of the last fit's streamed pass (0: the input was not folded)."""
        return int(self._lib.dr4g_last_fold(self._h))

    @property
    def last_proj_path(self) -> str:
        return {0: "none", 1: "cuda_cores", 2: "tcgen05", 3: "dmma_f64"}[int(self._lib.dr4g_last_proj_path(self._h))]

    @property
    def launch_count(self) -> int:
        return int(self._lib.dr4g_launch_count(self._h))
