#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
timeout 600 python -m pytest tests/test_gpu_gram_tc.py tests/test_gpu_parity.py tests/test_gpu_streaming.py -x -q -m gpu 2>&1 | tail -3
timeout 120 python tools/gram_bench.py ragged 2 "500000,516;1000000,784;1000000,1000;1000000,300;1000000,132" 2>&1 | grep "^ragged"
