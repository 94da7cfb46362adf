#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
DR4G_TRI_TRACE=1 timeout 120 python tools/eigh_probe2.py "256,32" 2>&1 | grep -A5 "tridiag trace" | head -14
DR4G_TRI_TRACE=1 timeout 120 python tools/eigh_probe2.py "512,32" 2>&1 | grep -A5 "tridiag trace" | head -7
DR4G_TRI_TRACE=1 timeout 120 python tools/eigh_probe2.py "128,32" 2>&1 | grep -A5 "tridiag trace" | head -7
