#!/bin/bash
# ncu --set full of the pipelined K3 kernel on the small trace workload (one launch)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 120 python tools/trace_x3.py 0 1 > gpurun_out/r2f_plain.log 2>&1 && \
timeout -s KILL 280 ncu --set full --clock-control none --import-source on -k regex:mlp_x3p_kernel -s 1 -c 1 -f -o gpurun_out/r2f_x3p python tools/trace_x3.py 0 1 > gpurun_out/r2f_ncu.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/r2f_ncu.log; head -1 gpurun_out/r2f_plain.log
