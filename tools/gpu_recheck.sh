#!/bin/bash
# re-entry check: GPU suite (with durations), smoke, default bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
T0=$(date +%s); timeout 2400 python -m pytest tests -q -m gpu -x --durations=15 > gpurun_out/recheck_pytest.log 2>&1; echo "pytest rc=$? elapsed $(( $(date +%s) - T0 )) s"; tail -25 gpurun_out/recheck_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
T0=$(date +%s); python bench.py > gpurun_out/recheck_bench.json 2> gpurun_out/recheck_bench.err; echo "bench rc=$? elapsed $(( $(date +%s) - T0 )) s"
python - <<'PY'
import json
j = json.loads(open('gpurun_out/recheck_bench.json').read().strip().splitlines()[-1])
print('C4', j['value'], j['ms_per_step'], j['stage_ms'], j['parity_check'], j['clocks'], 'frac', j['roofline']['frac'])
print('e2e', j['e2e']); print('cpu', j['cpu_baseline']); print('launches', j['gpu_launches'])
for k, v in (j.get('also') or {}).items():
    print(k, round(v.get('ms_per_step', 0), 3), v.get('gram_kernel'), {a: round(b, 3) for a, b in (v.get('stage_ms') or {}).items()}, 'T', (v.get('transform') or {}).get('ms'), 'roof', (v.get('roofline') or {}).get('frac'), (v.get('parity_check') or {}).get('sigma_rel'), v.get('error'))
PY
