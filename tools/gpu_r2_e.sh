#!/bin/bash
# CTA-0 timeline of K3 under a debug-flag set (usage: gpu_r2_e.sh FLAGS)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
X3_DBG=${1:-0} timeout -s KILL 60 python tools/trace_x3.py 0 1200 > gpurun_out/r2e_trace.txt 2>&1; head -1 gpurun_out/r2e_trace.txt
