#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
timeout 600 python -m pytest tests/test_gpu_project_tc.py tests/test_gpu_project_f64.py tests/test_gpu_parity.py tests/test_gpu_streaming.py -x -q -m gpu 2>&1 | tail -4
bash tools/gpu_shapes.sh 2>&1 | grep -E "d=   50|d=   99|d= 1001|k= 200|d=  254"
