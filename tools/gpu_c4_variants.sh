#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
v() { local tag=$1; shift; env "$@" timeout 600 python bench.py --steps 8 --warmup 3 --also none --no-cpu-baseline --e2e-steps 0 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C4 $tag', round(j['ms_per_step'],2), 'gram', round(j['stage_ms']['gram'],2), 'eig', round(j['stage_ms']['eigensolve'],2), j['clocks']['sm_mhz'], 'sig', j['parity_check']['sigma_rel'])"; }
v base
v merge DR4G_MERGE_DIAG=1
v merge_nopace DR4G_MERGE_DIAG=1 DR4G_PACE_MB=0
v nopace DR4G_PACE_MB=0
v merge_pace96 DR4G_MERGE_DIAG=1 DR4G_PACE_MB=96
v base2
