#!/bin/bash
# which part of the pair Gram's pipeline sets its pace: timings with the MMAs, the conversion, or both switched off
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gram_tc.py -x -q -m gpu 2>&1 | tail -2
run() { local tag=$1; shift; env "$@" timeout 300 python tools/gram_bench.py $tag ${SPLIT:-2} "4194304,512;2097152,1024;1000000,256" 2>&1 | grep "^$tag" | sed 's/err([0-9]* rows)=[^ ]* //'; }
run full
run full_op4 DR4G_OP_STAGES=4
run nomma DR4G_GRAM_TRACE=2
run noconv DR4G_GRAM_TRACE=4
run noconv_nomma DR4G_GRAM_TRACE=6
SPLIT=1 run tf32
DR4G_GRAM_TRACE=$((60*256+17)) timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/gt_off.txt 2>&1
sed -n 86,99p gpurun_out/gt_off.txt | cut -c1-100
DR4G_GRAM_TRACE=$((17)) timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/gt_diag.txt 2>&1
sed -n 86,99p gpurun_out/gt_diag.txt | cut -c1-100
