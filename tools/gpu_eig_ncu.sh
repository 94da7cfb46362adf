#!/bin/bash
# --set full capture of the cluster tridiagonalisation (d = 256, the C2 eigensolve) and the per-kernel times of the eigensolve
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 200 python tools/eigh_probe2.py "256,32" > gpurun_out/eig_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tridiag_cluster_kernel -s 2 -c 1 -f -o gpurun_out/r02_tridiag_d256 python tools/eigh_probe2.py "256,32" > gpurun_out/eig_ncu.log 2>&1
echo "capture rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02_eig_launches_d256.csv python tools/eigh_probe2.py "256,32" > /dev/null 2>&1
python - <<'PY'
import csv
rows = [r for r in csv.reader(open('gpurun_out/r02_eig_launches_d256.csv')) if len(r) > 10]
h = rows[0]
ki, vi = h.index('Kernel Name'), h.index('Metric Value')
for r in rows[1:14]:
    print(f"{r[ki][:80]:80s} {float(r[vi].replace(',',''))/1e3:10.1f} us  grid={r[h.index('Grid Size')]} block={r[h.index('Block Size')]}")
PY
