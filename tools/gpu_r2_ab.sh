#!/bin/bash
# same-box A/B of K3 debug-flag variants on a quarter of config 2 (usage: gpu_r2_ab.sh "0,4")
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
X3_SCALE=${2:-40} X3_AB="$1" timeout -s KILL 200 python tools/trace_x3.py 2>&1 | tail -12 | tee gpurun_out/r2_ab.txt
