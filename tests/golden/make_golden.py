"""Generate tests/golden/*.npz from the UNMODIFIED reference, run in the build container.

    make -C oracle
    python tests/golden/make_golden.py

The reference's own Python package (``/root/reference/dimreduce4gpu``) is imported and driven
through its public estimators with ``backend="cpu"`` and ``DIMREDUCE4GPU_CPU_LIB_PATH`` pointing
at ``oracle/_ref/libdimreduce4cpu.so`` (the reference CPU backend compiled as-is).  The inputs are
the reference's own test inputs (``tests/test_cpu_backend.py:27,49-52``,
``tests/test_sklearn_parity.py:23-24,50-51``, ``tests/test_sklearn_parity_comprehensive.py:29-62``)
plus BASELINE.json config 1 (10,000 x 100, k=10) and two mean-offset / spiked cases used by the
centring and power-iteration parity tests.  ``/root/reference`` does not exist on the GPU box,
so what travels is the committed ``.npz`` files; each stores the input (or, for config 1, the seed
and a checksum of the input), and every output array of the reference estimator.
"""
from __future__ import annotations

import os
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REPO = HERE.parent.parent
REF = Path("/root/reference")


def main() -> None:
    so = REPO / "oracle" / "_ref" / "libdimreduce4cpu.so"
    if not so.is_file() or not REF.is_dir():
        sys.exit("needs /root/reference and oracle/_ref (make -C oracle)")
    os.environ["DIMREDUCE4GPU_CPU_LIB_PATH"] = str(so)
    sys.path.insert(0, str(REF))
    sys.path.insert(0, str(HERE))
    import dimreduce4gpu as ref  # the reference package itself
    import inputs

    assert Path(ref.__file__).resolve().is_relative_to(REF)

    for name, (kind, X, k, alg) in inputs.cases().items():
        est = (ref.PCA if kind == "pca" else ref.TruncatedSVD)(
            n_components=k, algorithm=alg, n_iter=5, random_state=42, backend="cpu")
        Z = est.fit_transform(X)
        out = dict(
            components=est.components_, singular_values=est.singular_values_,
            explained_variance=est.explained_variance_, explained_variance_ratio=est.explained_variance_ratio_,
            x_sha256=inputs.sha(X), x_shape=np.asarray(X.shape), k=np.int32(k), kind=kind, algorithm=alg,
        )
        if getattr(est, "mean_", None) is not None:
            out["mean"] = est.mean_
        if X.size <= 1 << 16:  # small: keep every output of the reference in full
            out["U"] = est._U
            out["X_transformed"] = Z
        else:  # large: head rows + column norms of the scores
            out["Xt_head"] = Z[:64].copy()
            out["U_head"] = est._U[:64].copy()
            out["Xt_colnorm"] = np.linalg.norm(Z.astype(np.float64), axis=0)
        np.savez_compressed(HERE / f"{name}.npz", **out)
        print(f"{name}: X{X.shape} k={k} sigma0={out['singular_values'][0]:.6g}")


if __name__ == "__main__":
    main()
