#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
run() { local tag=$1; shift; env "$@" timeout 300 python tools/gram_bench.py $tag ${SPLIT:-2} "${SHAPES:-4194304,512;2097152,1024;1000000,768}" 2>&1 | grep "^$tag" | sed 's/err([0-9]* rows)=[^ ]* //; s/path=[^ ]* //'; }
run base
for p in 4 12 32 64; do run pace$p DR4G_PACE_MB=$p; done
timeout 600 env DR4G_PACE_MB=16 python -m pytest tests/test_gpu_gram_tc.py -x -q -m gpu 2>&1 | tail -2
