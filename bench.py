#!/usr/bin/env python
"""bench.py - PCA/TruncatedSVD fit throughput (rows/s) on B200, BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c2|c3|c1|c5|c5s|r1|r2|n1] [--impl reference]

A step is one fit (the whole hot path: pilot shift, streamed Gram + column statistics, exchange between ranks, centring,
eigensolve, outputs) over one device-resident row shard of synthetic data.
  c4 (default, every N): PCA k=32, 2^25 x 512 fp32 per GPU = 256M x 512 over 8 GPUs    (BASELINE.json configs[3], the
                         north-star target; 64 GiB per GPU, so BENCH is the 1-GPU point of the weak-scaling curve)
  c2: PCA k=32, 1,000,000 x 256 fp32 per GPU                                          (configs[1])
  c3: TruncatedSVD algorithm='power' k=64, 4,194,304 x 1024 fp32                       (configs[2])
  c1: TruncatedSVD k=10 (default algorithm 'power'), 10,000 x 100 fp32                 (configs[0])
  c5: PCA k=100, 2,500,000 x 4096 fp64 per GPU = 20M x 4096 over 8 GPUs                (configs[4]); c5s: 2^18 rows of it
  r1, r2: the reference's own bench shapes (bench/run_benchmarks.py:45-53): 10,000 x 128 k=32 PCA, 50,000 x 256 k=64 TSVD power
  n1: PCA k=8, 2^24 x 32 fp32: a narrow input (HBM-bound, AI = 16 flop/B), streamed with its rows folded 8-fold
At N = 1 the secondary workloads (c2, c3, c1, r1, r2, n1, c5s) run short in the same process and are attached under
"also".  N > 1: one process per GPU (torchrun), rows sharded, one NCCL collective on the d x d partials per fit, weak
scaling (fixed shard per GPU).  Prints ONE JSON line on rank 0.
`--impl reference` times the reference's own CPU backend (oracle/_ref/libdimreduce4cpu.so, else the numpy
restatement) on a bounded row sample of the same workload on the host cores."""
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
    "c1": dict(kind="tsvd", n=10_000, d=100, k=10, algorithm="power", dtype="f32",
               name="TruncatedSVD k=10 10000x100 fp32 (configs[0])"),
    "c2": dict(kind="pca", n=1_000_000, d=256, k=32, algorithm="cusolver", dtype="f32",
               name="PCA k=32 1Mx256 fp32 per GPU (configs[1])"),
    "c3": dict(kind="tsvd", n=4 * 2**20, d=1024, k=64, algorithm="power", dtype="f32",
               name="TruncatedSVD power k=64 4Mx1024 fp32 (configs[2])"),
    "c4": dict(kind="pca", n=2**25, d=512, k=32, algorithm="cusolver", dtype="f32",
               name="PCA k=32 2^25x512 fp32 per GPU (configs[3] shard: 256Mx512 over 8 GPUs)"),
    "c5": dict(kind="pca", n=2_500_000, d=4096, k=100, algorithm="cusolver", dtype="f64",
               name="PCA k=100 2.5Mx4096 fp64 per GPU (configs[4] shard: 20Mx4096 over 8 GPUs)"),
    "c5s": dict(kind="pca", n=2**18, d=4096, k=100, algorithm="cusolver", dtype="f64",
                name="PCA k=100 262144x4096 fp64 (configs[4] family, reduced n)"),
    "r1": dict(kind="pca", n=10_000, d=128, k=32, algorithm="cusolver", dtype="f32",
               name="PCA cusolver k=32 10000x128 fp32 (reference bench/run_benchmarks.py:50)"),
    "r2": dict(kind="tsvd", n=50_000, d=256, k=64, algorithm="power", dtype="f32",
               name="TruncatedSVD power k=64 50000x256 fp32 (reference bench/run_benchmarks.py:51)"),
    "n1": dict(kind="pca", n=2**24, d=32, k=8, algorithm="cusolver", dtype="f32",
               name="PCA k=8 2^24x32 fp32 (narrow input: rows folded 8-fold into the 256-column pass)"),
}
ALSO_DEFAULT = ("c2", "c3", "c1", "r1", "r2", "n1", "c5s")


def _peaks() -> dict:
    """Roofline denominators: the driver's MEASURED_PEAKS.json (HBM copy, bf16 dense) plus this repo's own probe of the
    TF32 and FP64 dense GEMM rates on the same pool (tools/measure_peaks.py -> profiles/r02_measured_peaks.json)."""
    out = dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback (B200_PROFILING.md)")
    p = ROOT / "MEASURED_PEAKS.json"
    if p.is_file():
        j = json.loads(p.read_text())
        out.update(hbm_gbs=float(j["hbm_gbs"]), bf16_tflops=float(j["bf16_tflops"]),
                   bf16_tflops_sustained=float(j.get("bf16_tflops_sustained", j["bf16_tflops"])), source="measured")
    out.update(tf32_tflops=out["bf16_tflops"] / 2.0, tf32_tflops_sustained=out["bf16_tflops_sustained"] / 2.0,
               fp64_tflops=37.0, fp64_tflops_sustained=37.0, tensor_source="proxy: bf16 / 2 (TF32), nominal 37 TF/s (FP64)")
    q = ROOT / "profiles" / "r02_measured_peaks.json"
    if q.is_file():
        j = json.loads(q.read_text())
        out.update(tf32_tflops=float(j["tf32_tflops"]), tf32_tflops_sustained=float(j["tf32_tflops_sustained"]),
                   fp64_tflops=float(j["fp64_tflops"]), fp64_tflops_sustained=float(j["fp64_tflops_sustained"]),
                   tensor_source="measured (tools/measure_peaks.py, profiles/r02_measured_peaks.json)")
    return out


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
    dt = torch.float32 if wl["dtype"] == "f32" else torch.float64
    g = torch.Generator(device=device)
    g.manual_seed(1234 + rank)
    X = torch.empty(n, d, dtype=dt, device=device)
    scale = torch.linspace(1.0, 3.0, d, device=device, dtype=dt)
    off = torch.linspace(-5.0, 5.0, d, device=device, dtype=dt) if wl["kind"] == "pca" else None
    slab = max(1, (256 << 20) // (X.element_size() * d))
    for r0 in range(0, n, slab):
        v = X[r0:r0 + slab]
        v.normal_(generator=g)
        v.mul_(scale)
        if off is not None:
            v.add_(off)
    return X


def cpu_reference_time(wl: dict, rows: int, repeats: int, warmup: int):
    """The reference CPU backend (or its restatement) on the first `rows` rows of rank 0's recipe.  The reference has
    float entry points only (include/cpu_backend.h:32-51): an fp64 workload is timed through them in fp32."""
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


def _cpu_rows(wl: dict, budget: float) -> int:
    """Rows the CPU backend gets per call: cost ~ n d^2 (sgesdd / randomized passes), sized from `budget` rows at d=256."""
    return int(min(wl["n"], max(5_000, budget * (256.0 / wl["d"]) ** 2)))


def run_reference(args, wl: dict) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warm = args.steps, args.warmup
    rows = _cpu_rows(wl, 3_000_000 / max(1, steps + warm))
    kind, cores, ts = cpu_reference_time(wl, rows, steps, warm)
    total = float(sum(ts))
    value = rows * steps / total
    sample = (f"first {rows} rows x {wl['d']} of the workload per step ({kind} CPU backend"
              f"{', fp32 entry points' if wl['dtype'] == 'f64' else ''}, OpenBLAS threads={cores}); rows/s is linear in n")
    line = dict(metric="fit_rows_per_s", value=value, unit="rows/s", n_gpus=args.gpus, steps=steps, warmup=warm,
                ms_per_step=1e3 * total / steps, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="f32", data="synthetic", impl="reference",
                config=dict(workload=wl["name"], rows_per_step=rows, n_features=wl["d"], n_components=wl["k"],
                            algorithm=wl["algorithm"]),
                cpu_baseline=dict(value=value, unit="rows/s", cores=cores, kind=kind, sample=sample),
                e2e=dict(value=value, unit="rows/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
def issued_factor(path: str, d: int) -> float:
    """Tensor-pipe flops the Gram kernel issues per algorithmic flop (2 n d^2), in TF32-equivalents (a bf16
    instruction moves twice the multiplies per cycle, so a bf16 product counts 0.5)."""
    T = (d + 127) // 128
    S = (T + 1) // 2
    if path.startswith("tcgen05_pair"):
        # CTA-pair kernel: 256-wide super-tiles, upper block triangle; three products per off-diagonal
        # super-tile, two per diagonal one (G = P + R + R^T)
        f = (2.0 * S + 3.0 * (S * (S - 1) // 2)) * 65536.0 / (d * d)
        return f * (0.5 if path.endswith("bf16x3") else 1.0)
    if path == "tcgen05":
        tiles = 0
        for i in range(S):
            for j in range(i, S):
                na, nb = min(2, T - 2 * i), min(2, T - 2 * j)
                tiles += na * (na + 1) // 2 if i == j else na * nb
        return 3.0 * tiles / (T * T)
    return 0.0


def roofline(wl: dict, path: str, stages: dict, ms_per_step: float, peaks: dict, workload_key: str, fold: int = 0) -> dict | None:
    """Roofline of the dominant streamed kernel (the Gram pass) and of the whole fit.  Algorithmic work per row
    (SURVEY 8d / DESIGN.md 3): d*sizeof(T) bytes and 2*d^2 flops; one launch processes the whole shard.  Which bound
    binds is decided from the time each peak would need: HBM = bytes / measured copy bandwidth, tensor = flops the kernel
    has to issue / measured dense peak of the pipe it runs on."""
    gram_ms = stages.get("gram")
    if not gram_ms:
        return None
    n, d = wl["n"], wl["d"]
    f64 = wl["dtype"] == "f64"
    elem = 8 if f64 else 4
    bytes_alg, flops_alg = n * d * float(elem), 2.0 * n * d * d
    sec = gram_ms * 1e-3
    long_kernel = gram_ms > 5.0       # a kernel that runs for many ms sits at the power-capped clock
    if f64:
        peak_tf = peaks["fp64_tflops_sustained" if long_kernel else "fp64_tflops"]
        peak_name = "FP64 dense (DMMA)"
        issued = 0.5 * (1.0 + 128.0 / d) if d >= 128 else 1.0   # upper block triangle by 128
    else:
        peak_tf = peaks["tf32_tflops_sustained" if long_kernel else "tf32_tflops"]
        peak_name = "TF32 dense"
        # a narrow input folded f-fold runs the d' = f d kernel on n / f rows: f times the issued work per algorithmic flop
        issued = issued_factor(path, d * fold) * fold if fold else issued_factor(path, d)
    t_hbm = bytes_alg / (peaks["hbm_gbs"] * 1e9)
    t_tensor = issued * flops_alg / (peak_tf * 1e12) if issued else 0.0
    ach_tf, ach_gb = flops_alg / sec / 1e12, bytes_alg / sec / 1e9
    # (folded narrow inputs: the regime is the algorithmic one, AI = d/2 flop/B far below the ridge - the tensor work
    # the fold wastes on products of different rows is this implementation's choice and shows in tensor_pipe_frac)
    if t_tensor > t_hbm and not fold:
        roof = dict(bound="tensor", achieved=ach_tf, peak=peak_tf, unit="TFLOP/s", frac=ach_tf / peak_tf)
    else:
        roof = dict(bound="hbm", achieved=ach_gb, peak=peaks["hbm_gbs"], unit="GB/s", frac=ach_gb / peaks["hbm_gbs"])
    roof["traffic"] = None
    for name in ("r02_ncu_traffic.json", "r01_ncu_traffic.json"):
        tj = ROOT / "profiles" / name
        if tj.is_file():
            t = json.loads(tj.read_text()).get(workload_key)
            if t and t.get("gram_kernel", path) == path:
                roof["traffic"] = t["traffic_bytes_per_launch"]
                roof["traffic_source"] = t["source"]
                break
    step_s = ms_per_step * 1e-3
    dominant = max((k for k in stages if k != "total"), key=lambda k: stages[k])
    roof.update(kernel=f"gram ({path})", kernel_ms=gram_ms, peak_is=f"{peak_name}, {'sustained' if long_kernel else 'burst'}",
                peak_source=peaks["tensor_source"] if roof["bound"] == "tensor" else peaks["source"],
                arithmetic_intensity_flop_per_byte=d * 2.0 / elem,
                issued_over_algorithmic=issued, tensor_pipe_frac=issued * ach_tf / peak_tf if issued else None,
                achieved_tflops_algorithmic=ach_tf, achieved_gbs_algorithmic=ach_gb, hbm_frac=ach_gb / peaks["hbm_gbs"],
                t_hbm_ms=1e3 * t_hbm, t_tensor_ms=1e3 * t_tensor,
                fit_frac=(flops_alg / step_s / 1e12 / peak_tf) if roof["bound"] == "tensor"
                else (bytes_alg / step_s / 1e9 / peaks["hbm_gbs"]),
                dominant_stage=dominant, dominant_stage_ms=stages[dominant])
    if fold:
        roof["row_fold"] = fold
    if not f64 and path.endswith("bf16x3"):
        b = peaks["bf16_tflops_sustained" if long_kernel else "bf16_tflops"]
        roof["frac_of_bf16_peak"] = ach_tf / b
        roof["note"] = ("operands split into two bf16 terms, three kind::f16 products: frac is algorithmic fp32-GEMM flops "
                        "against the TF32 dense peak (the single-pass tensor rate of an fp32 input); tensor_pipe_frac counts "
                        "issued TF32-equivalents")
    return roof


def principal_angle(torch, A, B) -> float:
    """Largest principal angle (rad) between the row spaces of A and B (k x d, orthonormal rows)."""
    s = torch.linalg.svdvals(A.double() @ B.double().T)
    return float(torch.arccos(s.min().clamp(max=1.0)))


def parity_check(torch, dist, ctx, wl, X, out, world) -> dict:
    """Untimed, after the timed region: the timed fit (tcgen05 Gram) against the same fit with the generic fp64 DMMA
    Gram (gram_path=1) on the very shard that was timed; for N > 1 also bit-equality of components_ across ranks."""
    k, center = wl["k"], wl["kind"] == "pca"
    res = dict()
    path = ctx.last_gram_path
    ref_out = None
    if wl["dtype"] == "f32" and path != "generic_dmma":
        ctx.set_option("gram_path", 1)
        t0 = time.perf_counter()
        ref_out = ctx.fit(X, k, center=center, algorithm=wl["algorithm"], n_iter=5, random_state=42, tol=1e-5)
        torch.cuda.synchronize()
        res["reference_fit"] = (f"same shard, same algorithm and seed, gram_path=1 (fp64 DMMA products instead of the tcgen05 "
                                f"split), {time.perf_counter() - t0:.2f} s")
        ctx.set_option("gram_path", 0)
        s, sr = out["singular_values"].double(), ref_out["singular_values"].double()
        res["sigma_rel"] = float(((s - sr).abs() / sr).max())
        res["max_angle"] = principal_angle(torch, out["components"], ref_out["components"])
        ev, evr = out["explained_variance_ratio"].double(), ref_out["explained_variance_ratio"].double()
        res["evr_rel"] = float(((ev - evr).abs() / evr).max())
        if wl["algorithm"] != "cusolver" and wl["d"] > 256:
            res["note"] = ("'power' (subspace iteration, n_iter=5) on both sides with the same seed: the check isolates the Gram "
                           "kernel; on a flat spectrum the iteration amplifies Gram differences in the trailing Ritz vectors")
        if center:
            res["mean_abs"] = float((out["mean"].double() - ref_out["mean"].double()).abs().max())
    else:
        res["skipped"] = "timed fit already uses the fp64 DMMA Gram"
    if world > 1:
        mine = out["components"].contiguous()
        all_c = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(all_c, mine)
        res["ranks_equal"] = bool(all(torch.equal(all_c[0], c) for c in all_c[1:]))
    return res


def time_fit(torch, dist, ctx, wl, X, steps, warmup, world, sampler=None):
    """W warm-up steps, K timed steps bracketed by barrier + synchronize, CUDA events on the launching stream; then the
    same steps replayed one by one to read the library's per-stage events."""
    k, center = wl["k"], wl["kind"] == "pca"
    out = None

    def step():
        nonlocal out
        out = ctx.fit(X, k, center=center, algorithm=wl["algorithm"], n_iter=5, random_state=42, tol=1e-5, out=out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    mark = 0
    if sampler is not None:
        sampler.wait_first()
        mark = sampler.mark()
    launches0 = ctx.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = ctx.launch_count - launches0
    t = torch.tensor([ms_total], dtype=torch.float64, device=X.device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    per = ms_total / max(steps, 1)
    stage_acc: dict = {}
    replay = max(min(steps, 10), min(200, int(200.0 / max(per, 0.05))))
    if per > 200:
        replay = 2
    rt = torch.tensor([replay], device=X.device)
    if world > 1:
        dist.broadcast(rt, src=0)   # same count on every rank: the step contains collectives
    for _ in range(int(rt.item())):
        step()
        torch.cuda.synchronize()
        st = ctx.stage_times()
        for kk, v in (st or {}).items():
            stage_acc.setdefault(kk, []).append(v)
    stages = {kk: float(np.mean(v)) for kk, v in stage_acc.items()}
    return dict(ms_total=ms_total, ms_per_step=per, launches=int(launches), stages=stages, out=out, mark=mark, step=step,
                fold=ctx.last_fold)


def run_e2e(torch, dist, ctx, wl, X, rank, world, dev, steps):
    """Same metric through the reference-facing call with HOST buffers, host<->device copies inside the timed region:
    pinned shard in, fit (sharded when N > 1: one packed exchange), projection, components / U / X_transformed
    back out.  N = 1: the legacy C symbol the reference's ctypes layer binds (pca_float / truncated_svd_float /
    *_double).  N > 1: dr4g_fit_transform_host_* - the same routine behind the legacy symbols, on this rank's context and
    communicator (the reference has no multi-device entry point).  Plus, at N = 1, `pageable`: the drop-in estimator
    call PCA(...).fit_transform(np.ndarray) with pageable numpy input and outputs, as a user of the reference makes it."""
    import ctypes

    from dimreduce4gpu_b200.lib_dimreduce4gpu import _load_pca_lib, _load_tsvd_lib, params

    n, d, k = wl["n"], wl["d"], wl["k"]
    if steps < 1:
        return None
    f64 = wl["dtype"] == "f64"
    tdt, ndt, elem = (torch.float64, np.float64, 8) if f64 else (torch.float32, np.float32, 4)
    ne = int(min(n, (4 << 30) // (elem * d)))  # cap the pinned host copy at 4 GiB per rank
    Xh = torch.empty(ne, d, dtype=tdt, pin_memory=True)
    Xh.copy_(X[:ne])
    torch.cuda.synchronize()
    center = wl["kind"] == "pca"
    pin = lambda *s: torch.zeros(*s, dtype=tdt, pin_memory=True)  # noqa: E731
    Q, w, U, Xt, ev, evr, mean = pin(k, d), pin(k), pin(ne, k), pin(ne, k), pin(k), pin(k), pin(d)
    p = params()
    p.X_n, p.X_m, p.k = ne, d, k
    p.algorithm = wl["algorithm"].encode()
    p.n_iter, p.random_state, p.tol, p.verbose, p.gpu_id, p.whiten = 5, 42, 1e-5, 0, dev.index, False
    ts = []
    h2d, d2h = ne * d * elem, (2 * ne * k + k * d + 3 * k + (d if center else 0)) * elem
    if world == 1:
        ct = ctypes.c_double if f64 else ctypes.c_float
        fp = lambda t: ctypes.cast(t.data_ptr(), ctypes.POINTER(ct))  # noqa: E731
        sfx = "double" if f64 else "float"
        fn = _load_pca_lib(sfx) if center else _load_tsvd_lib(sfx)
        a = [fp(Xh), fp(Q), fp(w), fp(U), fp(Xt), fp(ev), fp(evr)] + ([fp(mean)] if center else []) + [p]
        for i in range(steps + 1):
            t0 = time.perf_counter()
            fn(*a)  # synchronous: returns after the D2H copies completed
            if i:
                ts.append(time.perf_counter() - t0)
        api = ("pca_" if center else "truncated_svd_") + sfx
    else:
        outs = dict(components=Q.numpy(), singular_values=w.numpy(), explained_variance=ev.numpy(),
                    explained_variance_ratio=evr.numpy(), mean=mean.numpy() if center else None,
                    X_transformed=Xt.numpy(), U=U.numpy())
        Xn = Xh.numpy()
        for i in range(steps + 1):
            dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ctx.fit_transform_host(Xn, k, center=center, algorithm=wl["algorithm"], n_iter=5, random_state=42, tol=1e-5,
                                   out=outs)
            if i:
                ts.append(time.perf_counter() - t0)
        api = "dr4g_fit_transform_host_" + ("double" if f64 else "float") + " (pinned host shard per rank, sharded fit)"
    t = torch.tensor([float(np.mean(ts))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res = dict(value=ne * world / float(t.item()), unit="rows/s", h2d_bytes_per_step=int(h2d),
               d2h_bytes_per_step=int(d2h), rows_per_gpu=ne, api=api, ms_per_step=1e3 * float(t.item()),
               host_memory="pinned")
    if world == 1:
        # the drop-in call with pageable numpy buffers (what truncated_svd.py:72-107 / pca.py:49-86 hand the library)
        import dimreduce4gpu_b200 as pkg

        npg = int(min(ne, (1 << 30) // (elem * d)))
        Xp = np.array(Xh[:npg].numpy(), copy=True)
        cls = pkg.PCA if center else pkg.TruncatedSVD
        est = cls(n_components=k, algorithm=wl["algorithm"], n_iter=5, random_state=42, backend="gpu",
                  dtype="float64" if f64 else None)
        tp = []
        for i in range(3):
            t0 = time.perf_counter()
            est.fit_transform(Xp)
            if i:
                tp.append(time.perf_counter() - t0)
        res["pageable"] = dict(value=npg / float(np.mean(tp)), unit="rows/s", rows=npg, ms_per_step=1e3 * float(np.mean(tp)),
                               api=f"{cls.__name__}(...).fit_transform(np.ndarray): pageable input, fresh np.zeros outputs")
    return res


def run_also(torch, dist, ctx, keys, rank, world, dev, peaks) -> dict:
    """Secondary BASELINE configs, short (3 warm-up + 10 steps each; c5s: 1 + 2), same timing rules."""
    res = {}
    for key in keys:
        wl = dict(WORKLOADS[key])
        try:
            X = make_shard(torch, wl, rank, dev)
            heavy = wl["dtype"] == "f64"
            r = time_fit(torch, dist, ctx, wl, X, 2 if heavy else 10, 1 if heavy else 3, world)
            path = ctx.last_gram_path
            entry = dict(workload=wl["name"], value=wl["n"] * world * (2 if heavy else 10) / (r["ms_total"] * 1e-3),
                         unit="rows/s", ms_per_step=r["ms_per_step"], steps=2 if heavy else 10, stage_ms=r["stages"],
                         gram_kernel=path, gpu_launches=r["launches"],
                         roofline=roofline(wl, path, r["stages"], r["ms_per_step"], peaks, key, r.get("fold", 0)))
            if wl["dtype"] == "f32":
                entry["parity_check"] = parity_check(torch, dist, ctx, wl, X, r["out"], world)
            # transform (K7) on the same shard: Z = (X - mean) Q^T, device resident
            o = r["out"]
            Z = ctx.transform(X, o["components"], mean=o["mean"])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 2 if heavy else 5
            e0.record()
            for _ in range(reps):
                ctx.transform(X, o["components"], mean=o["mean"], out=Z)
            e1.record()
            torch.cuda.synchronize()
            tms = e0.elapsed_time(e1) / reps
            tb = wl["n"] * (wl["d"] + wl["k"]) * X.element_size()
            entry["transform"] = dict(ms=tms, rows_per_s=wl["n"] / (tms * 1e-3), achieved_gbs=tb / (tms * 1e-3) / 1e9,
                                      hbm_frac=tb / (tms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                                      algorithmic_bytes=tb, kernel=ctx.last_proj_path)
            if heavy:   # fp64: AI ~ k/4 flop/B puts the projection on the fp64 pipe, not on HBM
                tf = 2.0 * wl["n"] * wl["d"] * wl["k"] / (tms * 1e-3) / 1e12
                entry["transform"].update(bound="tensor (FP64 DMMA)", achieved_tflops=tf, frac=tf / peaks["fp64_tflops"])
            res[key] = entry
            del X, Z, r, o
        except Exception as e:  # noqa: BLE001 - a secondary workload must not take the headline line down
            res[key] = dict(workload=wl["name"], error=f"{type(e).__name__}: {e}")
        torch.cuda.empty_cache()
    return res


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c4")
    ap.add_argument("--impl", choices=["native", "reference"], default="native")
    ap.add_argument("--gram-path", type=int, default=0, help="0 auto, 1 generic DMMA, 2 tcgen05")
    ap.add_argument("--gram-split", type=int, default=0, help="pair kernel: 0 auto (bf16x3), 1 3xTF32, 2 bf16x3")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--also", default="auto", help="comma list of secondary workloads, 'auto' (N=1: c2,c3,c1,r1,r2,n1,c5s), 'none'")
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
    if args.gram_split:
        ctx.set_option("gram_split", args.gram_split)

    n, d, k = wl["n"], wl["d"], wl["k"]
    X = make_shard(torch, wl, rank, dev)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    r = time_fit(torch, dist, ctx, wl, X, args.steps, args.warmup, world, sampler if rank == 0 else None)
    # a short timed region is shorter than nvidia-smi's sampling period: keep the same steps running (untimed)
    # until the sampler has seen the load at least 5 times, at most 1 s
    t_end = time.time() + 1.0
    while rank == 0 and world == 1 and sampler.proc is not None and sampler.mark() - r["mark"] < 5 and time.time() < t_end:
        r["step"]()
        torch.cuda.synchronize()
    clocks = sampler.stop(r["mark"]) if rank == 0 else None
    if clocks is not None:
        clocks["window"] = "timed steps + the same steps replayed untimed (stage events)"
    value = n * world * args.steps / (r["ms_total"] * 1e-3)
    path = ctx.last_gram_path
    stages = r["stages"]

    # ---- what was timed, checked (untimed)
    pc = parity_check(torch, dist, ctx, wl, X, r["out"], world)

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timed region
    e2e = run_e2e(torch, dist, ctx, wl, X, rank, world, dev, args.e2e_steps)

    peaks = _peaks()
    also = None
    keys = () if args.also == "none" else (tuple(x for x in ALSO_DEFAULT if x != args.workload) if args.also == "auto"
                                           else tuple(x for x in args.also.split(",") if x))
    if keys and (world == 1 or args.also != "auto"):
        del X
        r["out"] = None
        torch.cuda.empty_cache()
        also = run_also(torch, dist, ctx, keys, rank, world, dev, peaks)

    if rank == 0:
        roof = roofline(wl, path, stages, r["ms_per_step"], peaks, args.workload, r.get("fold", 0))
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            rows = _cpu_rows(wl, 400_000)
            kind, cores, ts = cpu_reference_time(wl, rows, 2, 1)
            v = rows / float(np.median(ts))
            cpu = dict(value=v, unit="rows/s", cores=cores, kind=kind,
                       sample=f"first {rows} rows x {d} of the workload, median of 2 after 1 warm-up"
                              f"{' (fp32 entry points: the reference CPU backend has no double ones)' if wl['dtype'] == 'f64' else ''}")
        elem = 8 if wl["dtype"] == "f64" else 4
        line = dict(metric="fit_rows_per_s", value=value, unit="rows/s", n_gpus=world, steps=args.steps,
                    warmup=args.warmup, ms_per_step=r["ms_per_step"], higher_is_better=True, scaling="weak",
                    vs_baseline=None, dtype=wl["dtype"], data="synthetic",
                    config=dict(workload=wl["name"], rows_per_gpu=n, n_features=d, n_components=k,
                                algorithm=wl["algorithm"], parallelism=f"row-shard x{world}",
                                l2_policy=f"input ({n * d * elem / 2**30:.1f} GiB per GPU) larger than the 126 MB L2"
                                if n * d * elem > (256 << 20) else "input smaller than L2: latency-bound config, L2-resident by design",
                                gram_kernel=path),
                    roofline=roof, cpu_baseline=cpu, e2e=e2e, gpu_launches=r["launches"], clocks=clocks,
                    stage_ms=stages, parity_check=pc, also=also)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
