#!/bin/bash
# GPU session 4 (= session 3 re-run after the container was replaced, plus the whole GPU suite): whole-stage TMA producers + cross-job pacing in the pair Gram, cp.async landing ring in the
# tcgen05 projection, small-d tridiagonalisation cluster sizes.
set -x
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gram_tc.py tests/test_gpu_project_tc.py tests/test_gpu_gram_f64.py -x -q -m gpu > gpurun_out/s3_pytest_kernels.log 2>&1; echo "kernels rc=$?"; tail -12 gpurun_out/s3_pytest_kernels.log
gram_probe() {
timeout 300 python - > gpurun_out/s3_gram_$1.log 2>&1 <<PY
import sys, torch
sys.path.insert(0, '.')
import dimreduce4gpu_b200 as pkg
ctx = pkg.DeviceContext(0)
ctx.set_option('gram_path', 2)
ctx.set_option('gram_split', $2)
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
    print(f"$1 split=$2 n={n} d={d} path={ctx.last_gram_path} ms={ms:.3f} alg_TF/s={2*n*d*d/ms/1e9:.1f} GB/s={n*d*4/ms/1e6:.0f}", flush=True)
    del X, G
PY
tail -4 gpurun_out/s3_gram_$1.log
}
gram_probe pace24 2
DR4G_PACE_MB=0 gram_probe pace0 2
DR4G_PACE_MB=8 gram_probe pace8 2
DR4G_PACE_MB=48 gram_probe pace48 2
gram_probe tf32_pace24 1
DR4G_OCOST_BF16=44 DR4G_OCOST2_BF16=70 gram_probe oc44_70 2
DR4G_GRAM_TRACE=$((60*256+1)) timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/s3_trace_off.txt 2>&1
DR4G_GRAM_TRACE=1 timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/s3_trace_diag.txt 2>&1
# projection timing
timeout 300 python - > gpurun_out/s3_proj.log 2>&1 <<'PY'
import sys, torch
sys.path.insert(0, '.')
import dimreduce4gpu_b200 as pkg
for path in (0, 1):
    ctx = pkg.DeviceContext(0); ctx.set_option('proj_path', path)
    for n, d, k in [(1000000,256,32),(4194304,512,32),(2097152,1024,64),(1000000,256,64),(500000,512,128)]:
        g = torch.Generator(device='cuda').manual_seed(n+d)
        X = torch.randn(n, d, device='cuda', generator=g)
        Q, _ = torch.linalg.qr(torch.randn(d, k, device='cuda', generator=g)); Q = Q.T.contiguous()
        mean = X[:4096].mean(0).contiguous()
        Z = ctx.transform(X, Q, mean=mean); torch.cuda.synchronize()
        ts = []
        for _ in range(7):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ctx.transform(X, Q, mean=mean, out=Z); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
        ms = sorted(ts)[3]
        print(f"proj path={ctx.last_proj_path} n={n} d={d} k={k} ms={ms:.3f} GB/s={n*(d+k)*4/ms/1e6:.0f}", flush=True)
        del X, Z
PY
cat gpurun_out/s3_proj.log
for cs in 1 2 4 8 16; do DR4G_TRI_CS=$cs timeout 120 python tools/eigh_probe2.py "64,8;100,10;128,32" 2>&1 | grep aniso | sed "s/^/cs=$cs /"; done > gpurun_out/s3_tri_cs.log; cat gpurun_out/s3_tri_cs.log
timeout 900 python bench.py --steps 5 --warmup 3 --also c2,c3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/s3_bench_c4.json 2> gpurun_out/s3_bench_c4.err; echo "bench rc=$?"; tail -c 600 gpurun_out/s3_bench_c4.err
python - <<'PY'
import json
j = json.loads(open('gpurun_out/s3_bench_c4.json').read().strip().splitlines()[-1])
print('C4', j['ms_per_step'], j['stage_ms'], j['parity_check'], j['clocks'])
for k, v in (j.get('also') or {}).items():
    print(k, v.get('ms_per_step'), v.get('stage_ms'), v.get('transform'), v.get('parity_check'), v.get('error'))
PY
timeout 300 python tools/ncu_target.py gram 4194304 512 > gpurun_out/s3_plain_g512.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gram_pair_kernel -s 1 -c 1 -o gpurun_out/r02_gram_pair_bf16_d512_paced python tools/ncu_target.py gram 4194304 512 > gpurun_out/s3_ncu_g512.log 2>&1
timeout 300 python tools/ncu_target.py project 1000000 256 32 > gpurun_out/s3_plain_proj.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:project_tc_kernel -s 1 -c 1 -o gpurun_out/r02_project_tc_c2_v2 python tools/ncu_target.py project 1000000 256 32 > gpurun_out/s3_ncu_proj.log 2>&1
ls -la gpurun_out/*.ncu-rep
cp gpurun_out/s3_bench_c4.json gpurun_out/s4_bench_c4.json
timeout 2400 python -m pytest tests -q -m gpu -x --durations=15 > gpurun_out/s4_pytest_all.log 2>&1; echo "all rc=$?"; tail -30 gpurun_out/s4_pytest_all.log
