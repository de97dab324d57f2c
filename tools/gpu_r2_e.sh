#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 60 python tools/trace_x3.py 0 1200 > gpurun_out/r2e_trace.txt 2>&1; head -1 gpurun_out/r2e_trace.txt
