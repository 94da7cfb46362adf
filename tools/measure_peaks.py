#!/usr/bin/env python
"""Probe: dense TF32 and FP64 GEMM peaks of this B200 (cuBLAS through torch.matmul - a PROBE, not the product:
BASELINE.md section 3 asks the builder to measure both before quoting a tensor-roofline fraction; the driver's
MEASURED_PEAKS.json holds only the bf16 figure and the HBM copy rate).

    python tools/measure_peaks.py [out.json]

For each dtype: best of 10 single GEMMs (burst) and the mean rate of GEMMs issued back to back for `--seconds`
(sustained, i.e. at the clock the 1 kW power cap settles on).  bf16 is re-measured the same way so that the ratio to
the driver's figure can be checked."""
from __future__ import annotations

import json
import subprocess
import sys
import time

import torch


def smi() -> str:
    try:
        return subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.sw_power_cap",
                               "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True, timeout=10).stdout.strip()
    except Exception as e:  # noqa: BLE001
        return f"unavailable: {e}"


def gemm_rate(dtype, n: int, seconds: float, tf32: bool = False) -> dict:
    torch.backends.cuda.matmul.allow_tf32 = tf32
    a = torch.randn(n, n, device="cuda", dtype=dtype)
    b = torch.randn(n, n, device="cuda", dtype=dtype)
    c = torch.empty(n, n, device="cuda", dtype=dtype)
    flops = 2.0 * n * n * n
    for _ in range(3):
        torch.matmul(a, b, out=c)
    torch.cuda.synchronize()
    best = 0.0
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, flops / (e0.elapsed_time(e1) * 1e-3))
    # sustained: back to back for `seconds`
    one = flops / best
    reps = max(3, int(seconds / one))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        torch.matmul(a, b, out=c)
    e1.record()
    mid = None
    t0 = time.time()
    while not e1.query():
        if mid is None and time.time() - t0 > seconds * 0.5:
            mid = smi()
        time.sleep(0.05)
    torch.cuda.synchronize()
    sus = flops * reps / (e0.elapsed_time(e1) * 1e-3)
    return dict(n=n, burst_tflops=best / 1e12, sustained_tflops=sus / 1e12, reps=reps, smi_mid_run=mid)


def main() -> None:
    out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/r02_measured_peaks.json"
    seconds = 3.0
    res = dict(gpu=torch.cuda.get_device_name(0), torch=torch.__version__, how=__doc__.split("\n\n")[2].strip(),
               smi_idle=smi())
    res["bf16"] = gemm_rate(torch.bfloat16, 8192, seconds)
    res["tf32"] = gemm_rate(torch.float32, 8192, seconds, tf32=True)
    res["fp32_simt"] = gemm_rate(torch.float32, 8192, seconds, tf32=False)
    res["fp64"] = gemm_rate(torch.float64, 8192, seconds)
    # HBM copy, same recipe as the driver (read + write bytes)
    x = torch.empty(1 << 30, dtype=torch.bfloat16, device="cuda")
    y = torch.empty_like(x)
    best = 0.0
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        y.copy_(x)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, 2 * x.numel() * 2 / (e0.elapsed_time(e1) * 1e-3))
    res["hbm_copy_gbs"] = best / 1e9
    res["tf32_tflops"] = res["tf32"]["burst_tflops"]
    res["tf32_tflops_sustained"] = res["tf32"]["sustained_tflops"]
    res["fp64_tflops"] = res["fp64"]["burst_tflops"]
    res["fp64_tflops_sustained"] = res["fp64"]["sustained_tflops"]
    with open(out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
