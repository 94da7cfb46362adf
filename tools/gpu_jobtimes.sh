#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
for mc in 48 52 56; do for shape in "4194304 512"; do
echo "== merge mcost $mc shape $shape"
DR4G_MERGE_DIAG=1 DR4G_MCOST_BF16=$mc DR4G_GRAM_TRACE=32 timeout 120 python tools/ncu_target.py gram $shape 2>&1 | grep -A3 "job times" | tail -3
done; done
for mc in 52; do
DR4G_MERGE_DIAG=1 DR4G_MCOST_BF16=$mc timeout 300 python tools/gram_bench.py mergeop8 2 "4194304,512;8388608,512" 2>&1 | grep "^merge" | sed 's/path=[^ ]* //'
DR4G_MERGE_DIAG=1 DR4G_MCOST_BF16=$mc DR4G_PACE_MB=0 timeout 300 python tools/gram_bench.py mergenopace 2 "4194304,512" 2>&1 | grep "^merge" | sed 's/path=[^ ]* //'
DR4G_MERGE_DIAG=1 DR4G_MCOST_BF16=$mc ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_bytes.sum --clock-control none -k regex:gram_pair_kernel -s 1 -c 1 python tools/ncu_target.py gram 4194304 512 2>&1 | grep -E "dram__|gpu__time|hit_rate|tensor|xbar2l1" | sed "s/^/merge /"
done
