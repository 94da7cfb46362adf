#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
export DR4G_OP_STAGES=${DR4G_OP_STAGES:-4}
timeout 300 python tools/ncu_target.py gram 4194304 512 > gpurun_out/ng_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gram_pair_kernel -s 1 -c 1 -f -o gpurun_out/r02_gram_pair_bf16_d512_v2 python tools/ncu_target.py gram 4194304 512 > gpurun_out/ng_ncu.log 2>&1
tail -2 gpurun_out/ng_ncu.log
timeout 900 python bench.py --steps 5 --warmup 3 --also c2,c3 --no-cpu-baseline --e2e-steps 0 > gpurun_out/ng_bench_c4.json 2> gpurun_out/ng_bench_c4.err; echo "bench rc=$?"; tail -c 300 gpurun_out/ng_bench_c4.err
python - <<'PY'
import json
j = json.loads(open('gpurun_out/ng_bench_c4.json').read().strip().splitlines()[-1])
print('C4', j['ms_per_step'], j['stage_ms'], j['parity_check'], j['clocks'], j['roofline']['frac'])
for k, v in (j.get('also') or {}).items():
    print(k, v.get('ms_per_step'), v.get('stage_ms'), v.get('transform', {}).get('ms'), v.get('parity_check', {}).get('sigma_rel'), v.get('error'))
PY
