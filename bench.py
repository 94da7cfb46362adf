#!/usr/bin/env python
"""bench.py - PCA/TruncatedSVD fit throughput (rows/s) on B200, BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4] [--impl reference]

A step is one fit (the whole hot path: pilot shift, streamed Gram + column statistics, allreduce,
centring, eigensolve, outputs) over one device-resident row shard of synthetic data.
  workload c2 (default): PCA k=32, 1,000,000 x 256 fp32 per GPU   (BASELINE.json configs[1])
  workload c3:           TruncatedSVD algorithm='power' k=64, 4,194,304 x 1024 fp32 (configs[2])
  workload c4:           PCA k=32, 2^25 x 512 fp32 per GPU = 256M x 512 over 8 GPUs (configs[3])
N > 1: one process per GPU (torchrun), rows sharded, one NCCL allreduce of the d x d partials per
fit, weak scaling (fixed shard per GPU).  Prints ONE JSON line on rank 0.
`--impl reference` times the reference's own CPU backend (oracle/_ref/libdimreduce4cpu.so, else the
numpy restatement) on a bounded row sample of the same workload on the host cores."""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    "c2": dict(kind="pca", n=1_000_000, d=256, k=32, algorithm="cusolver", name="PCA k=32 1Mx256 fp32 per GPU (configs[1])"),
    "c3": dict(kind="tsvd", n=4 * 2**20, d=1024, k=64, algorithm="power", name="TruncatedSVD power k=64 4Mx1024 fp32 (configs[2])"),
    "c4": dict(kind="pca", n=2**25, d=512, k=32, algorithm="cusolver", name="PCA k=32 2^25x512 fp32 per GPU (configs[3] shard)"),
}


def _peaks() -> dict:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.is_file():
        j = json.loads(p.read_text())
        return dict(hbm_gbs=float(j["hbm_gbs"]), bf16_tflops=float(j["bf16_tflops"]), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, source="fallback")  # B200_PROFILING.md fallback


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int) -> None:
        self.index, self.proc, self.lines = index, None, []

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self) -> None:
        for ln in self.proc.stdout:
            self.lines.append(ln)

    def wait_first(self, timeout: float = 5.0) -> None:
        """nvidia-smi needs a few hundred ms to come up: block until its first line so that a short timed
        region is not over before the first sample."""
        t0 = time.time()
        while self.proc is not None and not self.lines and time.time() - t0 < timeout:
            time.sleep(0.01)

    def mark(self) -> int:
        return len(self.lines)

    def stop(self, first: int = 0) -> dict:
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines[first:]:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def make_shard(torch, wl: dict, rank: int, device):
    """X = randn(n,d) * linspace(1,3,d) (+ linspace(-5,5,d) for PCA): the reference's own anisotropic
    test construction (tests/test_sklearn_parity_comprehensive.py:50-54) plus an offset so centring is
    exercised.  Generated on the device in slabs, seed = 1234 + rank."""
    n, d = wl["n"], wl["d"]
    g = torch.Generator(device=device)
    g.manual_seed(1234 + rank)
    X = torch.empty(n, d, dtype=torch.float32, device=device)
    scale = torch.linspace(1.0, 3.0, d, device=device)
    off = torch.linspace(-5.0, 5.0, d, device=device) if wl["kind"] == "pca" else None
    slab = max(1, (256 << 20) // (4 * d))
    for r0 in range(0, n, slab):
        v = X[r0:r0 + slab]
        v.normal_(generator=g)
        v.mul_(scale)
        if off is not None:
            v.add_(off)
    return X


def cpu_reference_time(wl: dict, rows: int, repeats: int, warmup: int):
    """The reference CPU backend (or its restatement) on the first `rows` rows of rank 0's recipe."""
    from oracle import cpu_oracle as O

    rng = np.random.default_rng(1234)
    d = wl["d"]
    X = rng.standard_normal((rows, d), dtype=np.float32) * np.linspace(1.0, 3.0, d, dtype=np.float32)
    if wl["kind"] == "pca":
        X += np.linspace(-5.0, 5.0, d, dtype=np.float32)
    cores = os.cpu_count() or 1
    if O.REF_SO.is_file():
        kind = "reference"
        lib = O.RefLib()
        O.RefLib.set_threads(cores)
        fn = (lambda: lib.pca(X, wl["k"], wl["algorithm"], 5, 42)) if wl["kind"] == "pca" else \
             (lambda: lib.tsvd(X, wl["k"], wl["algorithm"], 5, 42))
    else:
        kind = "port"
        fn = (lambda: O.restated_pca(X, wl["k"])) if wl["kind"] == "pca" else (lambda: O.restated_tsvd(X, wl["k"]))
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return kind, cores, ts


def run_reference(args, wl: dict) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warm = args.steps, args.warmup
    rows = int(min(wl["n"], max(20_000, 3_000_000 * (256.0 / wl["d"]) ** 2 // max(1, steps + warm))))
    kind, cores, ts = cpu_reference_time(wl, rows, steps, warm)
    total = float(sum(ts))
    value = rows * steps / total
    sample = f"first {rows} rows x {wl['d']} of the workload per step ({kind} CPU backend, OpenBLAS threads={cores})"
    line = dict(metric="fit_rows_per_s", value=value, unit="rows/s", n_gpus=args.gpus, steps=steps, warmup=warm,
                ms_per_step=1e3 * total / steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                data="synthetic", impl="reference",
                config=dict(workload=wl["name"], rows_per_step=rows, n_features=wl["d"], n_components=wl["k"],
                            algorithm=wl["algorithm"]),
                cpu_baseline=dict(value=value, unit="rows/s", cores=cores, kind=kind, sample=sample),
                e2e=dict(value=value, unit="rows/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


def run_e2e(torch, dist, ctx, wl, X, rank, world, dev, steps):
    """Same metric through the reference-facing call with HOST buffers.  N=1: the legacy C symbol the
    reference's ctypes layer binds (pca_float / truncated_svd_float: H2D of X, fit, transform, D2H of
    components, U and X_transformed).  N>1: per rank, pinned shard -> device, sharded fit, results -> host."""
    import ctypes

    from dimreduce4gpu_b200.lib_dimreduce4gpu import _load_pca_lib, _load_tsvd_lib, params

    n, d, k = wl["n"], wl["d"], wl["k"]
    ne = int(min(n, (4 << 30) // (4 * d)))  # cap the pinned host copy at 4 GiB per rank
    Xh = torch.empty(ne, d, dtype=torch.float32, pin_memory=True)
    Xh.copy_(X[:ne])
    torch.cuda.synchronize()
    center = wl["kind"] == "pca"
    if steps < 1:
        return None
    ts = []
    if world == 1:
        pin = lambda *s: torch.zeros(*s, dtype=torch.float32, pin_memory=True)  # noqa: E731
        Q, w, U, Xt, ev, evr, mean = pin(k, d), pin(k), pin(ne, k), pin(ne, k), pin(k), pin(k), pin(d)
        p = params()
        p.X_n, p.X_m, p.k = ne, d, k
        p.algorithm = wl["algorithm"].encode()
        p.n_iter, p.random_state, p.tol, p.verbose, p.gpu_id, p.whiten = 5, 42, 1e-5, 0, dev.index, False
        fp = lambda t: ctypes.cast(t.data_ptr(), ctypes.POINTER(ctypes.c_float))  # noqa: E731
        fn = _load_pca_lib() if center else _load_tsvd_lib()
        a = [fp(Xh), fp(Q), fp(w), fp(U), fp(Xt), fp(ev), fp(evr)] + ([fp(mean)] if center else []) + [p]
        for i in range(steps + 1):
            t0 = time.perf_counter()
            fn(*a)  # synchronous: returns after the D2H copies completed
            if i:
                ts.append(time.perf_counter() - t0)
        h2d, d2h = ne * d * 4, (2 * ne * k + k * d + 3 * k + (d if center else 0)) * 4
        api = "pca_float" if center else "truncated_svd_float"
    else:
        Xd = torch.empty(ne, d, dtype=torch.float32, device=dev)
        for i in range(steps + 1):
            dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            Xd.copy_(Xh, non_blocking=True)
            res = ctx.fit(Xd, k, center=center, algorithm=wl["algorithm"], n_iter=5, random_state=42, tol=1e-5)
            host = {kk: v.cpu() for kk, v in res.items() if v is not None}
            torch.cuda.synchronize()
            if i:
                ts.append(time.perf_counter() - t0)
        h2d, d2h = ne * d * 4, sum(v.numel() for v in host.values()) * 4
        api = "DeviceContext.fit on a pinned host shard per rank"
    t = torch.tensor([float(np.mean(ts))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return dict(value=ne * world / float(t.item()), unit="rows/s", h2d_bytes_per_step=int(h2d),
                d2h_bytes_per_step=int(d2h), rows_per_gpu=ne, api=api, ms_per_step=1e3 * float(t.item()))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c2")
    ap.add_argument("--impl", choices=["native", "reference"], default="native")
    ap.add_argument("--gram-path", type=int, default=0, help="0 auto, 1 generic DMMA, 2 tcgen05")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=3)
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if args.impl == "reference":
        run_reference(args, wl)
        return
    if args.warmup < 3:
        args.warmup = 3  # timing rule: W >= 3

    import torch
    import torch.distributed as dist

    from dimreduce4gpu_b200 import DeviceContext
    from dimreduce4gpu_b200._native import load_library

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product has no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    load_library()
    ctx = DeviceContext(local)
    if world > 1:
        ctx.init_comm_from_torch()
    if args.gram_path:
        ctx.set_option("gram_path", args.gram_path)

    n, d, k = wl["n"], wl["d"], wl["k"]
    center = wl["kind"] == "pca"
    X = make_shard(torch, wl, rank, dev)
    out = None

    def step():
        nonlocal out
        out = ctx.fit(X, k, center=center, algorithm=wl["algorithm"], n_iter=5, random_state=42, tol=1e-5, out=out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    if rank == 0:
        sampler.wait_first()
    mark = sampler.mark()
    launches0 = ctx.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = ctx.launch_count - launches0
    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    # per-kernel share: same steps again, synchronising each one to read the library's stage events; enough of
    # them (same count on every rank: the step contains collectives) for the clock sampler to see >= 0.2 s of load
    stage_acc: dict = {}
    replay = max(min(args.steps, 10), min(400, int(200.0 / max(ms_total / max(args.steps, 1), 0.05))))
    for _ in range(replay):
        step()
        torch.cuda.synchronize()
        st = ctx.stage_times()
        for kk, v in (st or {}).items():
            stage_acc.setdefault(kk, []).append(v)
    # a 30 ms timed region is shorter than nvidia-smi's sampling period: keep the same steps running (untimed)
    # until the sampler has seen the load at least 5 times, at most 1 s
    t_end = time.time() + 1.0
    while rank == 0 and world == 1 and sampler.proc is not None and sampler.mark() - mark < 5 and time.time() < t_end:
        step()
        torch.cuda.synchronize()
    clocks = sampler.stop(mark) if rank == 0 else None
    if clocks is not None:
        clocks["window"] = "timed steps + the same steps replayed untimed (stage events)"
    value = n * world * args.steps / (ms_total * 1e-3)

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timed region
    e2e = run_e2e(torch, dist, ctx, wl, X, rank, world, dev, args.e2e_steps)

    if rank == 0:
        peaks = _peaks()
        stages = {kk: float(np.mean(v)) for kk, v in stage_acc.items()}
        gram_ms = stages.get("gram")
        # Roofline of the dominant kernel (the streamed Gram pass).  Algorithmic work per row
        # (SURVEY 8d / DESIGN.md): d*4 bytes and 2*d^2 flops.  Which bound binds is decided from the
        # times each peak would need: HBM = bytes / measured copy bandwidth; tensor = tensor-pipe flops
        # the kernel must issue / TF32 dense peak (= measured bf16 / 2).  The tcgen05 kernel issues
        # 3 (3xTF32 split) x tiles_computed / tiles_total of the algorithmic flops; the generic kernel
        # runs fp64 DMMA and is reported against HBM.
        roof = None
        if gram_ms:
            tf32_peak = peaks["bf16_tflops"] / 2.0
            T = (d + 127) // 128
            S = (T + 1) // 2
            tiles = 0
            for I in range(S):
                for J in range(I, S):
                    na, nb = min(2, T - 2 * I), min(2, T - 2 * J)
                    tiles += na * (na + 1) // 2 if I == J else na * nb
            if ctx.last_gram_path == "tcgen05_pair":
                # CTA-pair kernel (csrc/gram_tc2.cu): 256-wide super-tiles, upper block triangle by 256;
                # three tf32 products per off-diagonal super-tile, two per diagonal one (G = P + R + R^T)
                issued_factor = (2.0 * S + 3.0 * (S * (S - 1) // 2)) * 65536.0 / (d * d)
            elif ctx.last_gram_path == "tcgen05":
                issued_factor = 3.0 * tiles / (T * T)
            else:
                issued_factor = 0.0
            bytes_alg, flops_alg = n * d * 4.0, 2.0 * n * d * d
            t_hbm = bytes_alg / (peaks["hbm_gbs"] * 1e9)
            t_tensor = issued_factor * flops_alg / (tf32_peak * 1e12)
            sec = gram_ms * 1e-3
            if t_tensor > t_hbm:
                ach = flops_alg / sec / 1e12
                roof = dict(bound="tensor", achieved=ach, peak=tf32_peak, unit="TFLOP/s", frac=ach / tf32_peak,
                            traffic=None, peak_is="TF32 dense = measured bf16 / 2",
                            issued_over_algorithmic=issued_factor,
                            tensor_pipe_frac=issued_factor * ach / tf32_peak)
            else:
                ach = bytes_alg / sec / 1e9
                roof = dict(bound="hbm", achieved=ach, peak=peaks["hbm_gbs"], unit="GB/s", frac=ach / peaks["hbm_gbs"],
                            traffic=None)
            # dram bytes of this kernel from the committed ncu capture of the same workload (per launch), if any
            tj = ROOT / "profiles" / "r01_ncu_traffic.json"
            if tj.is_file():
                t = json.loads(tj.read_text()).get(args.workload)
                if t and ctx.last_gram_path == "tcgen05_pair":
                    roof["traffic"] = t["traffic_bytes_per_launch"]
                    roof["traffic_source"] = t["source"]
            roof.update(kernel=f"gram ({ctx.last_gram_path})", kernel_ms=gram_ms, peak_source=peaks["source"],
                        arithmetic_intensity_flop_per_byte=d / 2.0,
                        achieved_tflops_algorithmic=flops_alg / sec / 1e12,
                        frac_of_3xtf32_effective_peak=flops_alg / sec / 1e12 / (tf32_peak / 3.0),
                        achieved_gbs_algorithmic=bytes_alg / sec / 1e9,
                        hbm_frac=bytes_alg / sec / 1e9 / peaks["hbm_gbs"])
        cpu = None
        if not args.no_cpu_baseline:
            rows = int(min(n, max(20_000, 400_000 * (256.0 / d) ** 2)))
            kind, cores, ts = cpu_reference_time(wl, rows, 2, 1)
            v = rows / float(np.median(ts))
            cpu = dict(value=v, unit="rows/s", cores=cores, kind=kind,
                       sample=f"first {rows} rows x {d} of the workload, median of 2 after 1 warm-up")
        line = dict(metric="fit_rows_per_s", value=value, unit="rows/s", n_gpus=world, steps=args.steps,
                    warmup=args.warmup, ms_per_step=ms_total / args.steps, higher_is_better=True, scaling="weak",
                    vs_baseline=None, dtype="f32", data="synthetic",
                    config=dict(workload=wl["name"], rows_per_gpu=n, n_features=d, n_components=k,
                                algorithm=wl["algorithm"], parallelism=f"row-shard x{world}",
                                l2_policy="input (>= 1 GB per GPU) larger than the 126 MB L2",
                                gram_kernel=ctx.last_gram_path),
                    roofline=roof, cpu_baseline=cpu, e2e=e2e, gpu_launches=int(launches), clocks=clocks,
                    stage_ms=stages)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
