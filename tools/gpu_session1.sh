#!/bin/bash
# GPU session 1 of round 2: new tests first (fast fail), peaks probe, Gram probes for both splits, bench lines.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/s1_smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_gram_tc.py -x -q -m gpu -s > gpurun_out/s1_pytest_gram.log 2>&1; echo "gram rc=$?"
tail -5 gpurun_out/s1_pytest_gram.log
timeout 600 python -m pytest tests/test_gpu_gram_f64.py -x -q -m gpu > gpurun_out/s1_pytest_f64.log 2>&1; echo "f64 gram rc=$?"
tail -5 gpurun_out/s1_pytest_f64.log
timeout 600 python -m pytest tests/test_gpu_project_tc.py -x -q -m gpu -s > gpurun_out/s1_pytest_proj.log 2>&1; echo "proj rc=$?"
tail -25 gpurun_out/s1_pytest_proj.log
timeout 300 python tools/measure_peaks.py gpurun_out/r02_measured_peaks.json > gpurun_out/s1_peaks.log 2>&1; echo "peaks rc=$?"
for split in 1 2; do
  for ocost in 36; do
  DR4G_SPLIT=$split timeout 300 python - > gpurun_out/s1_gram_split$split.log 2>&1 <<PY
import sys, time, torch
sys.path.insert(0, '.')
import dimreduce4gpu_b200 as pkg
ctx = pkg.DeviceContext(0)
ctx.set_option('gram_path', 2); ctx.set_option('gram_split', $split)
for n, d in [(1000000,256),(4194304,512),(2097152,1024),(1000000,768),(262144,2048)]:
    g = torch.Generator(device='cuda').manual_seed(n*7+d)
    X = torch.randn(n, d, device='cuda', generator=g) * torch.linspace(1,3,d,device='cuda') + 0.5
    shift = X[:4096].mean(0).contiguous()
    G = ctx.gram(X, shift); ctx.sync()
    ns = min(n, 200000)
    Xd = X[:ns].double() - shift.double()
    Gs = ctx.gram(X[:ns].contiguous(), shift); ctx.sync()
    Gr = Xd.T @ Xd
    dg = torch.sqrt(torch.diag(Gr))
    err = ((Gs - Gr).abs() / (dg[:,None]*dg[None,:])).max().item()
    ts = []
    for _ in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctx.gram(X, shift); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[3]
    print(f"split=$split n={n} d={d} path={ctx.last_gram_path} err(200k rows)={err:.2e} ms={ms:.3f} alg_TF/s={2*n*d*d/ms/1e9:.1f} GB/s={n*d*4/ms/1e6:.0f}", flush=True)
    del X, G
PY
  echo "split $split rc=$?"; cat gpurun_out/s1_gram_split$split.log | tail -8
  done
done
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/s1_bench_c4.json 2> gpurun_out/s1_bench_c4.err; echo "bench c4 rc=$?"
tail -c 3000 gpurun_out/s1_bench_c4.json; tail -5 gpurun_out/s1_bench_c4.err
timeout 600 python bench.py --steps 10 --warmup 3 --gram-split 1 --also none --no-cpu-baseline --e2e-steps 0 > gpurun_out/s1_bench_c4_tf32.json 2> gpurun_out/s1_bench_c4_tf32.err; echo "bench c4 tf32 rc=$?"
timeout 1500 python -m pytest tests/test_gpu_baseline_configs.py tests/test_gpu_eigh.py -x -q -m gpu -s > gpurun_out/s1_pytest_cfg.log 2>&1; echo "cfg rc=$?"
tail -15 gpurun_out/s1_pytest_cfg.log
