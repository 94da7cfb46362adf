"""Worker for the world_size > 1 tests (launched by torch.distributed.run).

mode=cpu : gloo.  Mirrors, in numpy, the exchange structure of fit_impl (csrc/api.cu, csrc/finalize.cu): rank-local
           pilot shifts, ONE all-gather of the packed [G | s | q | n | c_r] partials, re-basing to the common shift,
           replicated centring + eigensolve, and checks the result against the oracle on the concatenated rows.
mode=gpu : nccl, one process per GPU.  The real thing: DeviceContext.init_comm_from_torch + sharded
           dr4g_fit_float, compared with a single-GPU fit of the concatenated rows on rank 0 and
           with the oracle."""
from __future__ import annotations

import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def make_rows(rank: int, n: int, d: int) -> np.ndarray:
    rng = np.random.default_rng(4321 + rank)
    X = rng.standard_normal((n, d)) * np.linspace(1.0, 3.0, d) + np.linspace(-5.0, 5.0, d)
    return X.astype(np.float32)


def main() -> None:
    import torch
    import torch.distributed as dist

    from oracle import cpu_oracle as O

    mode, n, d, k = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    n_local = n + 7 * rank  # ragged shards
    X = make_rows(rank, n_local, d)

    if mode == "cpu":
        dist.init_process_group("gloo")
        Xd = X.astype(np.float64)
        rows = min(n_local, 4096)
        # every rank keeps its OWN pilot shift (held in fp32); ONE all-gather of [G | s | q | n | c_r]; every rank
        # re-bases the moments to the common shift c = sum n_r c_r / sum n_r and sums them in rank order
        c_r = (Xd[:rows].sum(0) / rows).astype(np.float32).astype(np.float64)
        Y = Xd - c_r
        mine = np.concatenate([(Y.T @ Y).ravel(), Y.sum(0), (Y * Y).sum(0), [float(n_local)], c_r])
        parts = [torch.zeros(mine.size, dtype=torch.float64) for _ in range(world)]
        dist.all_gather(parts, torch.from_numpy(mine))
        parts = [p.numpy() for p in parts]
        dd = d * d
        nt = sum(p[dd + 2 * d] for p in parts)
        c = sum(p[dd + 2 * d] * p[dd + 2 * d + 1:] for p in parts) / nt
        G, s, q = np.zeros((d, d)), np.zeros(d), np.zeros(d)
        for p in parts:
            Gr, sr, qr, nr, cr = p[:dd].reshape(d, d), p[dd:dd + d], p[dd + d:dd + 2 * d], p[dd + 2 * d], p[dd + 2 * d + 1:]
            dl = cr - c
            G += Gr + np.outer(sr, dl) + np.outer(dl, sr) + nr * np.outer(dl, dl)
            s += sr + nr * dl
            q += qr + 2.0 * dl * sr + nr * dl * dl
        C = G - np.outer(s, s) / nt
        mean = c + s / nt
        total_var = float(((q - s * s / nt) / (nt - 1)).sum())
        w, V = np.linalg.eigh(C)
        w, V = w[::-1][:k], V[:, ::-1][:, :k].T
        sv = np.sqrt(np.maximum(w, 0))
        ev = sv ** 2 / (nt - 1)
        gathered = [None] * world
        dist.all_gather_object(gathered, X)
        if rank == 0:
            ref = O.restated_pca(np.concatenate(gathered), k, dtype=np.float64)
            assert int(nt) == sum(g.shape[0] for g in gathered)
            np.testing.assert_allclose(sv, ref.singular_values, rtol=1e-9)
            np.testing.assert_allclose(ev, ref.explained_variance, rtol=1e-9)
            np.testing.assert_allclose(ev / total_var, ref.explained_variance_ratio, rtol=1e-9)
            np.testing.assert_allclose(mean, ref.mean, atol=1e-9)
            (Cr,), (Co,) = O.svd_flip_v(ref.components), O.svd_flip_v(V)
            assert O.max_principal_angle(Cr, Co) < 1e-6
            print("MULTIRANK_CPU_OK", flush=True)
        dist.barrier()
        dist.destroy_process_group()
        return

    import dimreduce4gpu_b200 as pkg

    def progress(msg):   # shows where a hang is when the launcher's timeout fires
        print(f"[rank {rank}] {msg}", file=sys.stderr, flush=True)

    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = pkg.DeviceContext(local)
    ctx.init_comm_from_torch()
    progress("communicator up")
    if mode == "gpu-fold":
        # narrow shards: every rank streams its rows folded (csrc/api.cu: gram_fold_factor), the exchange and the
        # re-basing see ordinary d x d partials; against the reference backend on the concatenated rows
        Xg = torch.from_numpy(X).cuda()
        res = ctx.fit(Xg, k, center=True, algorithm="cusolver")
        ctx.sync()
        assert ctx.last_fold == 256 // d, f"fold factor {ctx.last_fold}"
        mine = res["components"].clone()
        ref0 = mine.clone()
        dist.broadcast(ref0, src=0)
        assert torch.equal(mine, ref0), "ranks disagree on components_"
        gathered = [None] * world
        dist.all_gather_object(gathered, X)
        if rank == 0:
            Xall = np.concatenate(gathered)
            ref = O.RefLib().pca(Xall, k, "cusolver") if O.REF_SO.is_file() else O.restated_pca(Xall, k)
            s = res["singular_values"].cpu().numpy()
            err = float(np.max(np.abs(s - ref.singular_values) / ref.singular_values))
            ang = O.max_principal_angle(ref.components, res["components"].cpu().numpy())
            assert err < 1e-4 and ang < 1e-2, (err, ang)
            np.testing.assert_allclose(res["mean"].cpu().numpy(), ref.mean, atol=1e-4)
            print(f"MULTIRANK_FOLD_OK sigma {err:.1e} angle {ang:.1e} exchange {ctx.last_exchange_path}", flush=True)
        dist.barrier()
        dist.destroy_process_group()
        return
    Xg = torch.from_numpy(X).cuda()
    out = {}
    Z = None
    for alg in ("cusolver", "power"):
        res = ctx.fit(Xg, k, center=True, algorithm=alg, n_iter=60, random_state=3, tol=1e-9)
        ctx.sync()
        out[alg] = {kk: v.cpu().numpy() for kk, v in res.items() if v is not None}
        if alg == "cusolver":   # shard-local scores of the exact model (the flat spectrum leaves 'power' a rotated basis)
            Z = ctx.transform(Xg, res["components"], mean=res["mean"]).cpu().numpy()
        progress(f"sharded fit {alg} done")
    # the exchange ran as ONE kernel over NVLink peer memory (csrc/p2p.cu); NCCL's all-gather gives the same bits
    path = ctx.last_exchange_path
    progress(f"exchange path: {path}")
    if os.environ.get("DR4G_P2P_MAX_D", "1024") != "0":
        assert path == "peer_memory", f"exchange went through {path}"
        ctx.set_option("exchange", 1)
        res_n = ctx.fit(Xg, k, center=True, algorithm="cusolver")
        ctx.sync()
        assert ctx.last_exchange_path == "nccl"
        for name in ("components", "singular_values", "explained_variance_ratio", "mean"):
            assert np.array_equal(res_n[name].cpu().numpy(), out["cusolver"][name]), f"{name}: peer-memory and NCCL exchange differ"
        ctx.set_option("exchange", 0)
        for _ in range(5):   # repeated exchanges: both parities of the buffers, epochs in step on all ranks
            res_p = ctx.fit(Xg, k, center=True, algorithm="cusolver")
        ctx.sync()
        assert ctx.last_exchange_path == "peer_memory"
        assert np.array_equal(res_p["components"].cpu().numpy(), out["cusolver"]["components"])
        progress("peer-memory exchange checked against NCCL")
    # streaming fit across ranks: every rank feeds its shard in two chunks (the first call carries the pilot
    # allreduce), finalize all-reduces the packed accumulators: same model as the one-pass sharded fit
    progress("transform done")
    ctx.partial_fit_reset()
    half = n_local // 2
    ctx.partial_fit(Xg[:half].contiguous(), center=True)
    ctx.partial_fit(Xg[half:].contiguous(), center=True)
    part = ctx.partial_fit_finalize(d, k, True, torch.float32, algorithm="cusolver")
    ctx.sync()
    sp, so = part["singular_values"].cpu().numpy(), out["cusolver"]["singular_values"]
    assert np.max(np.abs(sp - so) / so) < 2e-5, "streaming sharded fit differs from the one-pass sharded fit"
    assert float((part["mean"].cpu() - torch.from_numpy(out["cusolver"]["mean"])).abs().max()) < 1e-5
    progress("streaming fit done")
    # every rank must hold the same model (replicated eigensolve on identical all-reduced inputs)
    mine = torch.from_numpy(out["cusolver"]["components"]).cuda()
    ref0 = mine.clone()
    dist.broadcast(ref0, src=0)
    assert torch.equal(mine, ref0), "ranks disagree on components_"
    gathered = [None] * world
    dist.all_gather_object(gathered, X)
    progress("gathered")
    if rank == 0:
        Xall = np.concatenate(gathered)
        ref = O.RefLib().pca(Xall, k, "cusolver") if O.REF_SO.is_file() else O.restated_pca(Xall, k)
        progress("oracle done")
        single = pkg.DeviceContext(local)
        one = single.fit(torch.from_numpy(Xall).cuda(), k, center=True)
        single.sync()
        s1 = one["singular_values"].cpu().numpy()
        for alg, tol in (("cusolver", 1e-4), ("power", 1e-3)):
            s = out[alg]["singular_values"]
            assert np.max(np.abs(s - ref.singular_values) / ref.singular_values) < tol, alg
        np.testing.assert_allclose(out["cusolver"]["singular_values"], s1, rtol=2e-5)
        np.testing.assert_allclose(out["cusolver"]["mean"], ref.mean, atol=1e-4)
        (Cr,), (Co,) = O.svd_flip_v(ref.components), O.svd_flip_v(out["cusolver"]["components"])
        assert O.max_principal_angle(Cr, Co) < 1e-2
        # shard-local transform: rank 0's rows of the reference scores
        Zr = (Xall[: X.shape[0]].astype(np.float64) - ref.mean) @ ref.components.T.astype(np.float64)
        sgn = np.sign(np.sum(Zr * Z, axis=0))
        assert np.max(np.abs(Zr - Z * sgn)) < 5e-3 * ref.singular_values[0] / np.sqrt(Xall.shape[0]) * 10
        print("MULTIRANK_GPU_OK", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    try:
        main()
    except BaseException:   # a failed rank must not leave its peer spinning in a collective until the launcher's timeout
        import traceback

        traceback.print_exc()
        sys.stderr.flush()
        os._exit(1)
