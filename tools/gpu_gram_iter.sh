#!/bin/bash
# one iteration of Gram-kernel work on the GPU: kernel tests, timings under a few env settings, pipeline traces
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gram_tc.py -x -q -m gpu > gpurun_out/gi_pytest.log 2>&1; echo "gram tests rc=$?"; tail -3 gpurun_out/gi_pytest.log
run() { local tag=$1; shift; env "$@" timeout 300 python tools/gram_bench.py $tag 2 > gpurun_out/gi_$tag.log 2>&1; grep "^$tag" gpurun_out/gi_$tag.log || tail -3 gpurun_out/gi_$tag.log; }
run op2 DR4G_OP_STAGES=2
run op4 DR4G_OP_STAGES=4
for v in "$@"; do run "$(echo $v | tr '= ' '__')" $v; done
DR4G_GRAM_TRACE=$((60*256+1)) timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/gi_trace_off.txt 2>&1
DR4G_GRAM_TRACE=$((61*256+1)) timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/gi_trace_off_r1.txt 2>&1
DR4G_GRAM_TRACE=1 timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/gi_trace_diag.txt 2>&1
sed -n 75,99p gpurun_out/gi_trace_off.txt | cut -c1-90
