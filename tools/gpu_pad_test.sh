#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_streaming.py tests/test_gpu_gram_f64.py -x -q -m gpu 2>&1 | tail -5
bash tools/gpu_shapes.sh 2>&1 | grep -E "d=   50|d=   99|d= 1001|d=   30|d=  254"
