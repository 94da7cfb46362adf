"""dimreduce4gpu_b200 - B200-native (sm_100a) PCA / TruncatedSVD fit+transform.

Drop-in for the reference package's public surface (dimreduce4gpu/__init__.py:1-66): same class
names, constructor arguments, methods and attributes; the native work happens in
lib/libdimreduce4gpu.so behind the C ABI in include/dimreduce4gpu.h."""
from __future__ import annotations

from ._backend import gpu_runnable, select_backend
from ._native import (get_library_path, native_available, native_built, native_runnable, require_native,
                      require_native_runnable)
from .lib_dimreduce4gpu import params
from .pca import PCA
from .truncated_svd import TruncatedSVD
from .device import DeviceContext

__version__ = "0.1.0"


def cpu_built() -> bool:
    """Reference API (lib_dimreduce4cpu.cpu_built): this build never ships a CPU backend."""
    return False


def require_cpu_built() -> None:
    raise RuntimeError("dimreduce4gpu_b200 is the CUDA (sm_100a) build; it has no CPU native library")


__all__ = ["PCA", "TruncatedSVD", "DeviceContext", "gpu_runnable", "native_built", "native_runnable",
           "native_available", "require_native", "require_native_runnable", "cpu_built", "require_cpu_built",
           "select_backend", "params", "get_library_path", "__version__"]
