#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for p in 0 8 32; do
  DR4G_PACE_MB=$p timeout 600 python bench.py --steps 8 --warmup 3 --also none --no-cpu-baseline --e2e-steps 0 > gpurun_out/pc_$p.json 2> gpurun_out/pc_$p.err
  python - $p <<'PY'
import json, sys
j = json.loads(open(f'gpurun_out/pc_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print('pace', sys.argv[1], 'ms', round(j['ms_per_step'], 2), 'gram', round(j['stage_ms']['gram'], 2), j['clocks']['sm_mhz'], j['clocks']['reasons'], 'sig', j['parity_check']['sigma_rel'])
PY
done
for p in 0 8 32; do
DR4G_PACE_MB=$p ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:gram_pair_kernel -s 1 -c 1 python tools/ncu_target.py gram 4194304 512 2>&1 | grep -E "dram__|gpu__time|hit_rate" | sed "s/^/pace$p d512 /"
DR4G_PACE_MB=$p ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:gram_pair_kernel -s 1 -c 1 python tools/ncu_target.py gram 2097152 1024 2>&1 | grep -E "dram__|gpu__time|hit_rate" | sed "s/^/pace$p d1024 /"
done
