"""PCA estimator - same surface as the reference (dimreduce4gpu/pca.py:15-94): TruncatedSVD of the
column-centred data, plus ``mean_``.  ``whiten=True`` is an opt-in the GPU applies (K6, csrc/postproc.cu): the rows
of ``components_`` are divided by sqrt(explained_variance_), so the scores have unit variance and
``inverse_transform`` undoes it.  That is what the reference's GPU code intends at src/tsvd.cu:195-207 (it indexes
lambda by feature there, a bug); the reference's CPU backend - the parity oracle - ignores the flag, and with
``whiten=False`` (default) nothing changes."""
from __future__ import annotations

from typing import Optional

from .truncated_svd import Backend, TruncatedSVD


class PCA(TruncatedSVD):
    """PCA on the B200 native backend (covariance method: one streamed pass over X)."""

    _center = True
    _entry = "pca"

    def __init__(self, n_components: int = 2, algorithm: str = "cusolver", n_iter: int = 5,
                 random_state: Optional[int] = None, tol: float = 1e-5, verbose: bool = False, gpu_id: int = 0,
                 whiten: bool = False, backend: Backend = "auto", *, dtype: Optional[str] = None) -> None:
        super().__init__(n_components=n_components, algorithm=algorithm, n_iter=n_iter, random_state=random_state,
                         tol=tol, verbose=verbose, gpu_id=gpu_id, backend=backend, dtype=dtype)
        self.whiten = bool(whiten)
