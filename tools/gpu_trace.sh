#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for t0 in 10 50; do
DR4G_GRAM_TRACE=$((t0*65536 + 60*256+17)) timeout 120 python tools/ncu_target.py gram 4194304 512 > gpurun_out/gt_off_$t0.txt 2>&1
DR4G_GRAM_TRACE=$((t0*65536 + 60*256+1)) timeout 120 python tools/ncu_target.py gram 4194304 512 > gpurun_out/gt_offp_$t0.txt 2>&1
echo "== t0=$t0 (converter stamps)"; sed -n 2p gpurun_out/gt_off_$t0.txt | cut -c1-110; sed -n 40,99p gpurun_out/gt_off_$t0.txt | cut -c1-100
echo "== t0=$t0 (producer stamps)"; sed -n 2p gpurun_out/gt_offp_$t0.txt | cut -c1-110; sed -n 70,99p gpurun_out/gt_offp_$t0.txt | cut -c1-100
done
