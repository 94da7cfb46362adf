#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
run() { local tag=$1; shift; env "$@" timeout 300 python tools/gram_bench.py $tag ${SPLIT:-2} "${SHAPES:-4194304,512;2097152,1024}" 2>&1 | grep "^$tag" | sed 's/err([0-9]* rows)=[^ ]* //; s/path=[^ ]* //'; }
run base
for oc in 28 32 40 44; do run oc$oc DR4G_OCOST_BF16=$oc; done
for oc2 in 52 68 76; do run oc2_$oc2 DR4G_OCOST2_BF16=$oc2; done
for p in 8 24 64; do run pace$p DR4G_PACE_MB=$p; done
