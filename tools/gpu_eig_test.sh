#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_eigh.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python tools/eigh_probe2.py "34,8;64,8;100,10;128,32;130,20;256,32" 2>&1 | grep "aniso\|lowrank"
