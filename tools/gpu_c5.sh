#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python bench.py --workload c5 --steps 2 --warmup 3 --also none --e2e-steps 1 > gpurun_out/c5_bench.json 2> gpurun_out/c5_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/c5_bench.err
python - <<'PY'
import json
j = json.loads(open('gpurun_out/c5_bench.json').read().strip().splitlines()[-1])
print('C5', j['value'], j['ms_per_step'], j['stage_ms'], j['parity_check'], j['clocks'], j['roofline'])
print('e2e', j['e2e']); print('cpu', j['cpu_baseline'])
PY
timeout 300 python tools/ncu_target.py gram64 65536 4096 > gpurun_out/c5_plain_g64.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gram_f64_kernel -s 1 -c 1 -f -o gpurun_out/r02_gram_f64_d4096 python tools/ncu_target.py gram64 65536 4096 > gpurun_out/c5_ncu_g64.log 2>&1
tail -2 gpurun_out/c5_ncu_g64.log
