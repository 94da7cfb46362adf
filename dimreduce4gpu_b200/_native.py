"""Locating and binding libdimreduce4gpu.so (the sm_100a build).

Mirrors the reference's discovery layer (dimreduce4gpu/_native.py:9-34 search order and the
``DIMREDUCE4GPU_LIB_PATH`` override, :73-120 driver probe, :136-200 friendly errors) with two
differences: the handle is dlopen()ed once and cached (the reference re-walks the paths and
re-``LoadLibrary``s on every fit, lib_dimreduce4gpu.py:48,91), and there is no CPU fallback -
every failure raises ``RuntimeError``.
"""
from __future__ import annotations

import ctypes
import os
import sys
import threading
from pathlib import Path

_LIBNAME = "libdimreduce4gpu.so"
_lock = threading.Lock()
_handle: ctypes.CDLL | None = None
_handle_path: str | None = None


def _candidate_paths() -> list[Path]:
    cands: list[Path] = []
    env = os.environ.get("DIMREDUCE4GPU_LIB_PATH")
    if env:
        p = Path(env).expanduser().resolve()
        cands.append(p.parent if p.is_file() else p)
    pkg = Path(__file__).resolve().parent
    cands.append(pkg / "lib")
    cands.append(Path(sys.prefix) / "dimreduce4gpu")
    cands.append(pkg.parent / "lib")
    return cands


_path_cache: dict[str | None, str] = {}   # DIMREDUCE4GPU_LIB_PATH value -> library found (successes only)


def get_library_path() -> str | None:
    """Full path of libdimreduce4gpu.so if it exists, else None (reference _native.py:37-50).  A hit is remembered per
    value of DIMREDUCE4GPU_LIB_PATH: the walk over the candidate directories (pathlib + lstat) cost ~25 us, four times
    per fit_transform."""
    key = os.environ.get("DIMREDUCE4GPU_LIB_PATH")
    hit = _path_cache.get(key)
    if hit is not None:
        return hit
    found = _find_library_path()
    if found is not None:
        _path_cache[key] = found
    return found


def _find_library_path() -> str | None:
    for base in _candidate_paths():
        try:
            c = base / _LIBNAME
            if c.is_file():
                return str(c)
        except OSError:
            continue
    return None


def native_library_path() -> Path | None:
    p = get_library_path()
    return Path(p) if p else None


def require_native_built() -> str:
    path = get_library_path()
    if path is None:
        raise RuntimeError(
            "dimreduce4gpu native CUDA library (libdimreduce4gpu.so) is not available. Build it with "
            "`python -m dimreduce4gpu_b200.build` (needs nvcc) or set DIMREDUCE4GPU_LIB_PATH. "
            f"Searched: {[str(p) for p in _candidate_paths()]}"
        )
    return path


def load_library() -> ctypes.CDLL:
    """dlopen once, bind every symbol of include/dimreduce4gpu.h, cache the handle."""
    global _handle, _handle_path
    if _handle is not None and _handle_path == get_library_path():   # fast path, no lock
        return _handle
    with _lock:
        path = require_native_built()
        if _handle is not None and _handle_path == path:
            return _handle
        try:
            lib = ctypes.CDLL(path)
        except OSError as e:
            raise RuntimeError(
                f"dimreduce4gpu found libdimreduce4gpu at {path} but the native CUDA library could not be loaded: {e}"
            ) from e
        from .lib_dimreduce4gpu import _bind  # late: avoids an import cycle

        _bind(lib)
        _handle, _handle_path = lib, path
        return lib


def native_built() -> bool:
    try:
        load_library()
        return True
    except RuntimeError:
        return False


native_available = native_built  # name used by the reference's examples/bench scripts


_driver_ok: tuple[bool, int, str] | None = None


def _cuda_driver_device_count() -> tuple[bool, int, str]:
    """(ok, count, reason) through the driver API, as the reference does (_native.py:73-120).  A positive answer is
    remembered for the life of the process (devices do not go away under a running process; a failure is re-probed)."""
    global _driver_ok
    if _driver_ok is not None:
        return _driver_ok
    res = _probe_cuda_driver()
    if res[0]:
        _driver_ok = res
    return res


def _probe_cuda_driver() -> tuple[bool, int, str]:
    try:
        libcuda = ctypes.CDLL("libcuda.so.1")
    except OSError as e:
        return False, 0, f"NVIDIA driver runtime missing (libcuda.so.1): {e}"
    cu_init, cu_count = libcuda.cuInit, libcuda.cuDeviceGetCount
    cu_init.argtypes, cu_init.restype = [ctypes.c_uint], ctypes.c_int
    cu_count.argtypes, cu_count.restype = [ctypes.POINTER(ctypes.c_int)], ctypes.c_int
    rc = int(cu_init(0))
    if rc != 0:
        return False, 0, f"cuInit failed with error code {rc}"
    n = ctypes.c_int(0)
    rc = int(cu_count(ctypes.byref(n)))
    if rc != 0:
        return False, 0, f"cuDeviceGetCount failed with error code {rc}"
    if n.value <= 0:
        return False, 0, "no CUDA devices detected"
    return True, int(n.value), ""


def native_runnable() -> bool:
    if not native_built():
        return False
    ok, _n, _why = _cuda_driver_device_count()
    return ok


def require_native_runnable() -> str:
    path = require_native_built()
    load_library()
    ok, _n, why = _cuda_driver_device_count()
    if not ok:
        raise RuntimeError(
            "dimreduce4gpu native CUDA library is present, but this environment cannot run GPU code "
            f"(this build has no CPU fallback). Reason: {why}"
        )
    return path


def require_native() -> str:
    return require_native_built()


def last_error() -> str:
    lib = load_library()
    msg = lib.dr4g_last_error()
    return msg.decode("utf-8", "replace") if msg else ""
