#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 300 python tools/eigh_probe2.py "4096,100;2048,64;1024,64;512,32;256,32;128,32" 2>&1 | grep "aniso\|flat"
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/eig_launches.csv python tools/eigh_probe2.py "4096,100;512,32;256,32" > /dev/null 2>&1
python - <<'PY'
import csv
rows = [r for r in csv.reader(open('gpurun_out/eig_launches.csv')) if len(r) > 10]
h = rows[0]
ki, vi = h.index('Kernel Name'), h.index('Metric Value')
for r in rows[1:]:
    name = r[ki]
    if 'dr4g' in name or 'kernel' in name:
        print(f"{name[:70]:70s} {float(r[vi].replace(',',''))/1e3:10.1f} us  grid={r[h.index('Grid Size')]} block={r[h.index('Block Size')]}")
PY
