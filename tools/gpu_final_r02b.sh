#!/bin/bash
# final evidence of round 2 (after row folding): smoke, default bench line + reference arm, launch list of the bench
# command, C4 launch traffic, --set full capture of the row-folded Gram pass (2^24 x 32 read as 2^21 x 256)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
T0=$(date +%s); python bench.py > gpurun_out/fin_bench.json 2> gpurun_out/fin_bench.err; echo "bench rc=$? elapsed $(( $(date +%s) - T0 )) s"
python - <<'PY'
import json
j = json.loads(open('gpurun_out/fin_bench.json').read().strip().splitlines()[-1])
print('C4', j['value'], j['ms_per_step'], j['stage_ms'], j['parity_check'], j['clocks'], 'frac', j['roofline']['frac'], 'traffic', j['roofline'].get('traffic'))
print('e2e', j['e2e']); print('cpu', j['cpu_baseline']); print('launches', j['gpu_launches'])
for k, v in (j.get('also') or {}).items():
    print(k, round(v.get('ms_per_step', 0), 3), v.get('gram_kernel'), {a: round(b, 3) for a, b in (v.get('stage_ms') or {}).items()}, 'T', (v.get('transform') or {}).get('ms'), (v.get('transform') or {}).get('hbm_frac'), 'roof', (v.get('roofline') or {}).get('bound'), (v.get('roofline') or {}).get('frac'), (v.get('parity_check') or {}).get('sigma_rel'), v.get('error'))
PY
T0=$(date +%s); python bench.py --impl reference > gpurun_out/fin_ref.json 2> gpurun_out/fin_ref.err; echo "ref rc=$? elapsed $(( $(date +%s) - T0 )) s"; tail -1 gpurun_out/fin_ref.json | cut -c1-400
B="python bench.py --steps 2 --warmup 3 --also none --no-cpu-baseline --e2e-steps 0"
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:dr4g -c 400 --csv --log-file gpurun_out/r02_launches_c4.csv $B > gpurun_out/fin_bench_ncu.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/r02_launches_c4.csv
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.avg.per_second,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:gram_pair_kernel -s 1 -c 1 --csv --log-file gpurun_out/r02_c4_gram_traffic.csv python tools/ncu_target.py gram 33554432 512 > gpurun_out/fin_ncu_c4gram.log 2>&1
echo "c4 traffic rc=$?"
timeout 300 python tools/ncu_target.py fit 16777216 32 8 > gpurun_out/fin_plain_fold.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gram_pair_kernel -s 1 -c 1 -f -o gpurun_out/r02_gram_folded_d32 python tools/ncu_target.py fit 16777216 32 8 > gpurun_out/fin_ncu_fold.log 2>&1
echo "fold capture rc=$?"; cat gpurun_out/fin_plain_fold.log | tail -1
timeout 300 python tools/ncu_target.py project 16777216 32 8 > gpurun_out/fin_plain_pfold.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:project_tc_kernel -s 1 -c 1 -f -o gpurun_out/r02_project_folded_d32 python tools/ncu_target.py project 16777216 32 8 > gpurun_out/fin_ncu_pfold.log 2>&1
echo "proj fold capture rc=$?"; cat gpurun_out/fin_plain_pfold.log | tail -1
