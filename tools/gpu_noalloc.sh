#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
echo "--- alone, fresh process"; timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -x 2>&1 | tail -4
echo "--- after the other tests"; timeout 2400 python -m pytest tests -q -m gpu -x 2>&1 | tail -4
