#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
cat > /tmp/small.py <<'PY'
import sys, torch
sys.path.insert(0, '.')
import dimreduce4gpu_b200 as pkg
ctx = pkg.DeviceContext(0)
n, d, k, center = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
X = torch.randn(n, d, device='cuda') * torch.linspace(1, 3, d, device='cuda')
for _ in range(3):
    out = ctx.fit(X, k, center=bool(center), algorithm='cusolver')
ctx.sync()
print(ctx.stage_times())
PY
for cfg in "10000 100 10 0" "10000 128 32 1" "1000000 256 32 1"; do
python /tmp/small.py $cfg
ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:dr4g -c 60 --csv --log-file gpurun_out/small.csv python /tmp/small.py $cfg > /dev/null 2>&1
python - "$cfg" <<'PY'
import csv, sys
rows=[r for r in csv.reader(open('gpurun_out/small.csv')) if len(r)>10]
h=rows[0]; ki,vi=h.index('Kernel Name'),h.index('Metric Value')
names=[(r[ki].split('(')[0].split('::')[-1][:40], float(r[vi].replace(',',''))/1e3, r[h.index('Grid Size')]) for r in rows[1:]]
per=len(names)//3
print(sys.argv[1], 'launches per fit', per)
for nm, us, g in names[-per:]: print(f"   {nm:42s} {us:8.1f} us grid={g}")
print('   sum', round(sum(u for _,u,_ in names[-per:]),1), 'us')
PY
done
