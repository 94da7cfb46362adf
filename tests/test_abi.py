"""C-ABI checks that need no GPU: the library loads, exports every symbol include/dimreduce4gpu.h
declares (the reference's CI contract is `nm -D` >= ci/expected_symbols.txt,
ci/verify_native_so.sh:57-78), `params` has the reference's layout, and the product refuses to run
without a GPU instead of falling back to a CPU path."""
from __future__ import annotations

import ctypes
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "dimreduce4gpu.h"


@pytest.fixture(scope="module")
def lib_path():
    import dimreduce4gpu_b200 as pkg

    p = pkg.get_library_path()
    if p is None:
        from dimreduce4gpu_b200.build import build_native

        p = str(build_native())
    return p


def _declared_symbols():
    text = HEADER.read_text()
    return sorted(set(re.findall(r"^DR4G_EXPORT [\w\s\*]+?\b(\w+)\(", text, flags=re.M)))


def test_header_declares_reference_contract_symbols():
    syms = _declared_symbols()
    # reference ci/expected_symbols.txt:4-5 + the double entry points of include/tsvd.h:49, include/pca.h:49
    for s in ("pca_float", "truncated_svd_float", "pca_double", "truncated_svd_double"):
        assert s in syms
    assert len(syms) >= 28


def test_library_exports_every_declared_symbol(lib_path):
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    missing = [s for s in _declared_symbols() if s not in exported]
    assert not missing, missing
    from dimreduce4gpu_b200.lib_dimreduce4gpu import EXPORTED_SYMBOLS

    assert set(EXPORTED_SYMBOLS) == set(_declared_symbols())


def test_library_loads_and_has_no_undefined_symbols(lib_path):
    lib = ctypes.CDLL(lib_path)
    lib.dr4g_abi_version.restype = ctypes.c_int
    assert lib.dr4g_abi_version() == 1
    r = subprocess.run(["ldd", "-r", lib_path], capture_output=True, text=True)
    assert "undefined symbol" not in r.stdout + r.stderr
    # no link-time dependency on the CUDA toolkit's shared libraries (cudart is static, NCCL is dlopen'ed)
    assert not re.search(r"libcublas|libcusolver|libcusparse|libnccl|libcudart", r.stdout)


def test_params_layout_matches_reference():
    from dimreduce4gpu_b200 import params

    # 48 bytes, X_n@0 X_m@4 k@8 algorithm@16 n_iter@24 random_state@28 tol@32 verbose@36 gpu_id@40 whiten@44
    assert ctypes.sizeof(params) == 48
    off = {n: getattr(params, n).offset for n, _ in params._fields_}
    assert off == dict(X_n=0, X_m=4, k=8, algorithm=16, n_iter=24, random_state=28, tol=32, verbose=36, gpu_id=40,
                       whiten=44)


def test_public_surface_matches_reference():
    import inspect

    import dimreduce4gpu_b200 as pkg

    sig = inspect.signature(pkg.PCA.__init__)
    assert list(sig.parameters)[1:10] == ["n_components", "algorithm", "n_iter", "random_state", "tol", "verbose",
                                          "gpu_id", "whiten", "backend"]
    assert sig.parameters["algorithm"].default == "cusolver"
    sig = inspect.signature(pkg.TruncatedSVD.__init__)
    assert list(sig.parameters)[1:9] == ["n_components", "algorithm", "n_iter", "random_state", "tol", "verbose",
                                         "gpu_id", "backend"]
    assert sig.parameters["algorithm"].default == "power"
    for m in ("fit", "fit_transform", "transform", "inverse_transform"):
        assert callable(getattr(pkg.PCA, m))
    est = pkg.PCA(3)
    with pytest.raises(AttributeError):
        est.components_
    with pytest.raises(AttributeError):
        est.singular_values_
    for name in ("native_built", "native_runnable", "require_native", "gpu_runnable", "select_backend", "cpu_built"):
        assert hasattr(pkg, name)


def test_no_cpu_fallback():
    import dimreduce4gpu_b200 as pkg

    X = np.arange(12, dtype=np.float32).reshape(4, 3)
    with pytest.raises(RuntimeError):
        pkg.PCA(2, backend="cpu").fit_transform(X)
    if not pkg.native_runnable():
        # reference tests/test_api_cpu.py:35-39: clear RuntimeError naming cuda/native/libdimreduce4gpu
        with pytest.raises(RuntimeError) as ei:
            pkg.PCA(2, backend="gpu").fit_transform(X)
        msg = str(ei.value).lower()
        assert "cuda" in msg or "native" in msg or "libdimreduce4gpu" in msg
        with pytest.raises(RuntimeError):
            pkg.TruncatedSVD(2, backend="auto").fit_transform(X)


def test_product_never_imports_oracle():
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|cpu_oracle|libdimreduce4cpu", re.M)
    for py in (ROOT / "dimreduce4gpu_b200").glob("*.py"):
        assert not pat.search(py.read_text()), py
    for cu in (ROOT / "dimreduce4gpu_b200" / "csrc").glob("*"):
        assert "libdimreduce4cpu" not in cu.read_text(), cu


def test_library_path_cache_follows_the_environment_variable(lib_path, monkeypatch, tmp_path):
    """get_library_path() remembers a hit per value of DIMREDUCE4GPU_LIB_PATH (the reference re-walks the directories on
    every call, _native.py:37-50): changing the variable must still be honoured, and a miss is not remembered."""
    from dimreduce4gpu_b200 import _native

    from pathlib import Path

    monkeypatch.delenv("DIMREDUCE4GPU_LIB_PATH", raising=False)
    default = _native.get_library_path()
    assert default == str(lib_path)
    assert _native.get_library_path() == default   # cached hit
    copy = tmp_path / "libdimreduce4gpu.so"
    copy.write_bytes(Path(lib_path).read_bytes())
    monkeypatch.setenv("DIMREDUCE4GPU_LIB_PATH", str(copy))
    assert _native.get_library_path() == str(copy)
    monkeypatch.setenv("DIMREDUCE4GPU_LIB_PATH", str(tmp_path / "nowhere"))
    assert _native.get_library_path() == default          # falls through to the package directory, like the reference
    monkeypatch.delenv("DIMREDUCE4GPU_LIB_PATH")
    assert _native.get_library_path() == default
