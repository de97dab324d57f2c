#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 200 python tools/ab_frame.py 2>&1 | tee gpurun_out/r2e_ab.txt | tail -12
