#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
DR4G_GRAM_TRACE=$((60*256+17)) timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/gt_off.txt 2>&1
DR4G_GRAM_TRACE=$((17)) timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/gt_diag.txt 2>&1
DR4G_GRAM_TRACE=$((60*256+1)) timeout 120 python tools/ncu_target.py gram 1048576 512 > gpurun_out/gt_off_p.txt 2>&1
sed -n 1,2p gpurun_out/gt_off.txt | cut -c1-110
sed -n 80,99p gpurun_out/gt_off.txt | cut -c1-100
sed -n 84,99p gpurun_out/gt_diag.txt | cut -c1-100
sed -n 1,2p gpurun_out/gt_off_p.txt | cut -c1-110
sed -n 84,99p gpurun_out/gt_off_p.txt | cut -c1-100
