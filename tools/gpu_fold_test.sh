#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_fold.py tests/test_gpu_reference_suite.py -q -m gpu -x 2>&1 | tail -15
timeout 300 python bench.py --workload n1 --also none --e2e-steps 1 > gpurun_out/fold_n1.json 2> gpurun_out/fold_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/fold_n1.err
python - <<'PY'
import json
j = json.loads(open('gpurun_out/fold_n1.json').read().strip().splitlines()[-1])
print('n1', j['value'], j['ms_per_step'], j['stage_ms'], j.get('parity_check'), 'roof', j['roofline'])
print('e2e', j['e2e'])
PY
