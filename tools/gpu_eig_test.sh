#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_eigh.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python tools/eigh_probe2.py "512,32;384,32;300,20;258,16;256,32;128,32;100,10;34,8;7,3" 2>&1 | grep "aniso"
DR4G_TRI_NOREG=1 timeout 300 python tools/eigh_probe2.py "512,32;256,32" 2>&1 | grep "aniso" | sed 's/^/noreg /'
DR4G_TRI_TRACE=1 timeout 120 python tools/eigh_probe2.py "512,32" 2>&1 | grep -A5 "tridiag trace" | head -6
DR4G_TRI_TRACE=1 timeout 120 python tools/eigh_probe2.py "256,32" 2>&1 | grep -A5 "tridiag trace" | head -6
