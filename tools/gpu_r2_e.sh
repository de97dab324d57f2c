#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for f in 0 32; do X3_DBG=$f timeout -s KILL 60 python tools/trace_x3.py 0 600 > gpurun_out/r2e_trace_$f.txt 2>&1; head -1 gpurun_out/r2e_trace_$f.txt; done
