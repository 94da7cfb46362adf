"""Random-shape sweep (tools/fuzz_shapes.py) as a test: fit + transform on shapes drawn around the dispatch boundaries - tiny
inputs, row folding and its tails, padded chunk copies, single-CTA / CTA-pair Gram, k slices - against fp64 references
formed on the GPU.  Gates: sigma 1e-4 (fp32) / 1e-10 (fp64), scores 2e-5 / 1e-11 of max|Z| (BASELINE.json north_star)."""
from __future__ import annotations

import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu

TOOL = Path(__file__).resolve().parents[1] / "tools" / "fuzz_shapes.py"


@pytest.mark.parametrize("seed,cases,dtype", [(11, 70, "f32"), (12, 70, "f32"), (13, 60, "f64")])
def test_random_shapes_around_the_dispatch_boundaries(seed, cases, dtype):
    r = subprocess.run([sys.executable, str(TOOL), str(seed), str(cases), dtype], capture_output=True, text=True, timeout=900)
    tail = "\n".join(ln for ln in r.stdout.splitlines() if not ln.startswith("ok "))
    assert r.returncode == 0 and "FUZZ PASSED" in r.stdout, tail[-3000:] + r.stderr[-2000:]


HOST_TOOL = Path(__file__).resolve().parents[1] / "tools" / "fuzz_host.py"


def test_random_shapes_through_the_host_pointer_symbols_around_the_staging_thresholds():
    """numpy in, numpy out (pca_* / truncated_svd_*, fp32 and fp64): inputs around 8 / 12 / 24 / 40 MB, score arrays around 4 / 8 MB
    - the sizes at which the pinned ring, the copy pool and the early page population switch on - against numpy fp64."""
    r = subprocess.run([sys.executable, str(HOST_TOOL), "5", "40"], capture_output=True, text=True, timeout=900)
    tail = "\n".join(ln for ln in r.stdout.splitlines() if not ln.startswith("ok "))
    assert r.returncode == 0 and "HOSTFUZZ PASSED" in r.stdout, tail[-3000:] + r.stderr[-2000:]
