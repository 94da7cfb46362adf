#!/bin/bash
# 8-GPU weak-scaling points of the north-star config (C4 shard per GPU) + C2 shards, as the driver launches them
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L | wc -l
for wl in c4 c2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --steps 10 --warmup 3 --workload $wl > gpurun_out/g8_bench_$wl.json 2> gpurun_out/g8_bench_$wl.err; echo "bench $wl rc=$?"
tail -c 300 gpurun_out/g8_bench_$wl.err | tail -3
python - $wl <<'PY'
import json, sys
j = json.loads(open(f'gpurun_out/g8_bench_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], 'value', j['value'], 'ms', j['ms_per_step'], j['stage_ms'], j['parity_check'], j['clocks'], 'e2e', j['e2e'])
PY
done
