#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T0=$(date +%s); python bench.py > gpurun_out/fb_bench.json 2> gpurun_out/fb_bench.err; echo "bench rc=$?"; echo "elapsed $(( $(date +%s) - T0 )) s"; tail -c 400 gpurun_out/fb_bench.err | grep -v "^\s" | tail -3
python - <<'PY'
import json
j = json.loads(open('gpurun_out/fb_bench.json').read().strip().splitlines()[-1])
print('C4', j['value'], j['ms_per_step'], j['stage_ms'], j['parity_check'], j['clocks'], j['roofline']['frac'], j['roofline']['peak'])
print('e2e', j['e2e']); print('cpu', j['cpu_baseline']); print('launches', j['gpu_launches'])
for k, v in (j.get('also') or {}).items():
    print(k, v.get('ms_per_step'), v.get('gram_kernel'), {a: round(b, 3) for a, b in (v.get('stage_ms') or {}).items()}, 'T', (v.get('transform') or {}).get('ms'), (v.get('transform') or {}).get('hbm_frac'), 'roof', (v.get('roofline') or {}).get('frac'), (v.get('roofline') or {}).get('bound'), (v.get('parity_check') or {}).get('sigma_rel'), v.get('error'))
PY
T0=$(date +%s); python bench.py --impl reference > gpurun_out/fb_ref.json 2> gpurun_out/fb_ref.err; echo "ref rc=$?"; echo "elapsed $(( $(date +%s) - T0 )) s"; cat gpurun_out/fb_ref.json | cut -c1-600
