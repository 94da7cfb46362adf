#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for p in 48 64 96; do
DR4G_PACE_MB=$p ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:gram_pair_kernel -s 1 -c 1 python tools/ncu_target.py gram 8388608 512 2>&1 | grep -E "dram__|gpu__time|hit_rate" | sed "s/^/pace$p n=2^23 d=512 /"
DR4G_PACE_MB=$p ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:gram_pair_kernel -s 1 -c 1 python tools/ncu_target.py gram 4194304 1024 2>&1 | grep -E "dram__|gpu__time|hit_rate" | sed "s/^/pace$p n=2^22 d=1024 /"
done
for p in 0 48 64 96; do DR4G_PACE_MB=$p timeout 300 python tools/gram_bench.py pace$p 2 "8388608,512;4194304,1024;1000000,768" 2>&1 | grep "^pace" | sed 's/err([0-9]* rows)=[^ ]* //; s/path=[^ ]* //'; done
for p in 0 64; do
  DR4G_PACE_MB=$p timeout 600 python bench.py --steps 8 --warmup 3 --also none --no-cpu-baseline --e2e-steps 0 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C4 pace $p', round(j['ms_per_step'],2), 'gram', round(j['stage_ms']['gram'],2), j['clocks']['sm_mhz'])"
done
