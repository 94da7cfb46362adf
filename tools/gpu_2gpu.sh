#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_multirank.py tests/test_gpu_streaming.py -q -m gpu -k "two_gpus or second_device" 2>&1 | tail -5
for wl in c4 c2; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 --workload $wl > gpurun_out/g2_bench_$wl.json 2> gpurun_out/g2_bench_$wl.err; echo "bench $wl rc=$?"
tail -c 300 gpurun_out/g2_bench_$wl.err
python - $wl <<'PY'
import json, sys
j = json.loads(open(f'gpurun_out/g2_bench_{sys.argv[1]}.json').read().strip().splitlines()[-1])
print(sys.argv[1], 'value', j['value'], 'ms', j['ms_per_step'], j['stage_ms'], j['parity_check'], j['clocks'], 'e2e', j['e2e'])
PY
done
