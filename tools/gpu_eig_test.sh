#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_eigh.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python tools/eigh_probe2.py "4096,100;2048,64;1024,64;1536,32" 2>&1 | grep "aniso"
