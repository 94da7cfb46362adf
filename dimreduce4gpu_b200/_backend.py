"""Backend selection, kept for API compatibility with the reference (dimreduce4gpu/_backend.py:45-52).

This build is GPU-only: "auto" and "gpu" both mean the sm_100a library; "cpu" is refused (the
reference's CPU backend is the parity oracle under oracle/, never a product path)."""
from __future__ import annotations

from typing import Literal

from ._native import native_runnable, require_native_runnable

Backend = Literal["auto", "gpu", "cpu"]


def gpu_runnable() -> bool:
    return native_runnable()


def select_backend(requested: Backend) -> Backend:
    if requested == "cpu":
        raise RuntimeError(
            "dimreduce4gpu_b200 is the CUDA (sm_100a) build and has no CPU backend or fallback; "
            "use backend='gpu' (or 'auto') on a machine with a B200 and libdimreduce4gpu.so built"
        )
    if requested not in ("auto", "gpu"):
        raise ValueError(f"unknown backend {requested!r}")
    require_native_runnable()  # raises RuntimeError naming cuda / libdimreduce4gpu when not runnable
    return "gpu"
