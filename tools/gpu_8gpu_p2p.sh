#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for p2p in 1024 0; do
DR4G_P2P_MAX_D=$p2p timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 8 --steps 50 --warmup 5 --workload c2 --e2e-steps 0 > gpurun_out/p2p8_$p2p.json 2> gpurun_out/p2p8_$p2p.err; echo "bench p2p=$p2p rc=$?"
python - $p2p <<'PY'
import json, sys
j = json.loads(open(f'gpurun_out/p2p8_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print('P2P_MAX_D', sys.argv[1], 'value', j['value'], 'ms', j['ms_per_step'], {k: round(v, 4) for k, v in j['stage_ms'].items()}, j['parity_check'].get('ranks_equal'))
PY
done
