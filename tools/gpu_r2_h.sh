#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 400 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_r2h.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r2h.log | cut -c1-300
timeout -s KILL 200 python tools/ab_frame.py 2>&1 | head -4
