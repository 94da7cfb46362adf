"""dr4g_eigh_topk_static_double (include/dimreduce4gpu.h): the exact path's eigensolver with its workspace in a
file-scope static device array - no cudaMalloc / cudaFree at call time (no
counterpart in the reference).  Checked through the C ABI against the existing entry point dr4g_eigh_topk_double on the
same matrices, twice in a row (the array keeps the last call's contents), at the gates of tests/test_gpu_eigh.py.
Collected last on purpose (zz): a failure here must not hide other tests under `pytest -x`."""
from __future__ import annotations

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _sym(d, seed):
    rng = np.random.default_rng(seed)
    X = rng.standard_normal((4 * d + 50, d)) * np.linspace(1, 3, d)
    return X.T @ X


def test_static_workspace_eigh_matches_the_arena_entry_point_twice_in_a_row():
    import torch

    import dimreduce4gpu_b200 as pkg
    from dimreduce4gpu_b200._native import load_library

    lib = load_library()
    assert lib.dr4g_eigh_topk_static_max_d() == 256
    ctx = pkg.DeviceContext(0)
    # two calls in a row with different sizes: the second finds the first one's data in the shared workspace
    for i, (d, k, seed) in enumerate([(256, 32, 1), (100, 100, 2), (256, 256, 3), (33, 7, 4)]):
        A = _sym(d, seed)
        Ad = torch.from_numpy(A).cuda()
        w0, V0 = ctx.eigh_topk(Ad, k)                      # existing entry point (arena workspace)
        ctx.sync()
        work = Ad.clone()
        w1 = torch.empty(k, dtype=torch.float64, device="cuda")
        V1 = torch.empty(k, d, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        ctx.eigh_topk_static(work, k, w1, V1)              # new entry point, straight through the C ABI
        ctx.sync()
        if i > 0:   # (the very first call of a process may still page in the module that holds the array)
            assert torch.cuda.mem_get_info()[0] == free0, "the static entry point changed the device's free memory"
        w0n, V0n, w1n, V1n = w0.cpu().numpy(), V0.cpu().numpy(), w1.cpu().numpy(), V1.cpu().numpy()
        ev = np.linalg.eigvalsh(A)
        scale = np.abs(ev).max()
        # same code on the same input: identical eigenvalues to rounding; vectors up to the gates of test_gpu_eigh.py
        assert np.max(np.abs(w1n - w0n)) / scale < 1e-13, (d, k)
        assert np.max(np.abs(w1n - ev[::-1][:k])) / scale < 1e-13, (d, k)
        assert np.max(np.linalg.norm(A @ V1n.T - V1n.T * w1n, axis=0)) / scale < 1e-12, (d, k)
        assert np.max(np.abs(V1n @ V1n.T - np.eye(k))) < 1e-11, (d, k)
        assert np.max(np.abs(np.abs(np.sum(V0n * V1n, axis=1)) - 1.0)) < 1e-9 or k == d, (d, k)


def test_static_workspace_eigh_rejects_sizes_beyond_its_maximum():
    import torch

    import dimreduce4gpu_b200 as pkg

    ctx = pkg.DeviceContext(0)
    d, k = 258, 4
    work = torch.eye(d, dtype=torch.float64, device="cuda")
    w = torch.empty(k, dtype=torch.float64, device="cuda")
    Vt = torch.empty(k, d, dtype=torch.float64, device="cuda")
    with pytest.raises(RuntimeError, match="beyond the static workspace"):
        ctx.eigh_topk_static(work, k, w, Vt)
