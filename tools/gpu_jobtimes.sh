#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
for pace in 0 8; do for shape in "4194304 512" "2097152 1024"; do
echo "== pace $pace shape $shape"
DR4G_PACE_MB=$pace DR4G_GRAM_TRACE=32 timeout 120 python tools/ncu_target.py gram $shape 2>&1 | grep -A12 "job times" | tail -14 | head -13
done; done
