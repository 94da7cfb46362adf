#!/bin/bash
# GPU session 2 of round 2: decoupled bf16x3 Gram, fp64 / projection / config tests, traces, ncu.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gram_tc.py -x -q -m gpu > gpurun_out/s2_pytest_gram.log 2>&1; echo "gram rc=$?"; tail -3 gpurun_out/s2_pytest_gram.log
gram_probe() {  # $1 = tag, env passed by caller
timeout 300 python - > gpurun_out/s2_gram_$1.log 2>&1 <<PY
import sys, torch
sys.path.insert(0, '.')
import dimreduce4gpu_b200 as pkg
ctx = pkg.DeviceContext(0)
ctx.set_option('gram_path', 2)
for n, d in [(1000000,256),(4194304,512),(2097152,1024),(1000000,768)]:
    g = torch.Generator(device='cuda').manual_seed(n*7+d)
    X = torch.randn(n, d, device='cuda', generator=g) * torch.linspace(1,3,d,device='cuda') + 0.5
    shift = X[:4096].mean(0).contiguous()
    G = ctx.gram(X, shift); ctx.sync()
    ts = []
    for _ in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctx.gram(X, shift); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[3]
    print(f"$1 n={n} d={d} path={ctx.last_gram_path} ms={ms:.3f} alg_TF/s={2*n*d*d/ms/1e9:.1f} GB/s={n*d*4/ms/1e6:.0f}", flush=True)
    del X, G
PY
cat gpurun_out/s2_gram_$1.log | tail -4
}
gram_probe base
DR4G_OCOST_BF16=30 DR4G_OCOST2_BF16=50 gram_probe oc30_50
DR4G_OCOST_BF16=44 DR4G_OCOST2_BF16=70 gram_probe oc44_70
DR4G_OCOST_BF16=52 DR4G_OCOST2_BF16=84 gram_probe oc52_84
# pipeline traces: a diagonal-job CTA (0) and an off-diagonal-job CTA (cluster 30 -> CTA 60) at d = 512
DR4G_GRAM_TRACE=1 timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/s2_trace_diag.txt 2>&1
DR4G_GRAM_TRACE=$((60*256+1)) timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/s2_trace_off.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_gram_f64.py tests/test_gpu_project_tc.py -q -m gpu > gpurun_out/s2_pytest_f64_proj.log 2>&1; echo "f64+proj rc=$?"; tail -15 gpurun_out/s2_pytest_f64_proj.log
timeout 900 python bench.py --steps 5 --warmup 3 --also c2,c5s --no-cpu-baseline --e2e-steps 0 > gpurun_out/s2_bench_c4.json 2> gpurun_out/s2_bench_c4.err; echo "bench rc=$?"; tail -c 600 gpurun_out/s2_bench_c4.err
python - <<'PY'
import json
j = json.loads(open('gpurun_out/s2_bench_c4.json').read().strip().splitlines()[-1])
print('C4', j['ms_per_step'], j['stage_ms'], j['parity_check'], j['clocks'])
for k, v in (j.get('also') or {}).items():
    print(k, v.get('ms_per_step'), v.get('stage_ms'), v.get('transform'), v.get('error'))
PY
timeout 2400 python -m pytest tests/test_gpu_baseline_configs.py tests/test_gpu_eigh.py tests/test_gpu_parity.py -q -m gpu > gpurun_out/s2_pytest_cfg.log 2>&1; echo "cfg rc=$?"; tail -25 gpurun_out/s2_pytest_cfg.log
# ncu: launch list of one C4-shaped fit family + full sets of the Gram at d = 512 / 1024 and the projection
timeout 300 python tools/ncu_target.py gram 4194304 512 > gpurun_out/s2_plain_g512.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gram_pair_kernel -s 1 -c 1 -o gpurun_out/r02_gram_pair_bf16_d512 python tools/ncu_target.py gram 4194304 512 > gpurun_out/s2_ncu_g512.log 2>&1
timeout 300 python tools/ncu_target.py gram 2097152 1024 > gpurun_out/s2_plain_g1024.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gram_pair_kernel -s 1 -c 1 -o gpurun_out/r02_gram_pair_bf16_d1024 python tools/ncu_target.py gram 2097152 1024 > gpurun_out/s2_ncu_g1024.log 2>&1
timeout 300 python tools/ncu_target.py project 1000000 256 32 > gpurun_out/s2_plain_proj.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:project_tc_kernel -s 1 -c 1 -o gpurun_out/r02_project_tc_c2 python tools/ncu_target.py project 1000000 256 32 > gpurun_out/s2_ncu_proj.log 2>&1
timeout 300 python tools/ncu_target.py gram64 65536 4096 > gpurun_out/s2_plain_g64.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gram_f64_kernel -s 1 -c 1 -o gpurun_out/r02_gram_f64_d4096 python tools/ncu_target.py gram64 65536 4096 > gpurun_out/s2_ncu_g64.log 2>&1
ls -la gpurun_out/*.ncu-rep
