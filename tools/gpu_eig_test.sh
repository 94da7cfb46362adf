#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_eigh.py -x -q -m gpu 2>&1 | tail -4
timeout 300 python tools/eigh_probe2.py "4096,100;1024,64;512,32;256,32;256,256;128,32" 2>&1 | grep "aniso\|flat\|lowrank"
timeout 600 python bench.py --workload c5s --steps 2 --warmup 3 --also none --e2e-steps 0 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('c5s', j['ms_per_step'], j['stage_ms'])"
