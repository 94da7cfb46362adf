#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
run() { local tag=$1; shift; env "$@" timeout 300 python tools/gram_bench.py $tag ${SPLIT:-2} "${SHAPES:-4194304,512;2097152,1024;1000000,768}" 2>&1 | grep "^$tag\|Error" | sed 's/path=[^ ]* //'; }
run base
for mc in 44 52 60 68; do run merge$mc DR4G_MERGE_DIAG=1 DR4G_MCOST_BF16=$mc; done
DR4G_MERGE_DIAG=1 timeout 600 python -m pytest tests/test_gpu_gram_tc.py -x -q -m gpu 2>&1 | tail -3
