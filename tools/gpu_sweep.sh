#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
run() { local tag=$1; shift; env "$@" timeout 300 python tools/gram_bench.py $tag ${SPLIT:-2} "${SHAPES:-4194304,512;2097152,1024;1000000,256}" 2>&1 | grep "^$tag" | sed 's/path=[^ ]* //'; }
for f in 256 512 1024 2048 4096 16384; do run flush$f FLUSH_ROWS=$f; done
