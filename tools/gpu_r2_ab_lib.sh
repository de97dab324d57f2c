#!/bin/bash
# same-box A/B of two builds of libucnerf.so (tools/ab/libucnerf_{old,new}.so) on the K3 microbenchmark
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for v in old new old new; do
  cp tools/ab/libucnerf_$v.so uc-nerf_b200/libucnerf.so
  echo -n "$v: "; X3_SCALE=40 X3_AB="0" timeout -s KILL 200 python tools/trace_x3.py 2>&1 | tail -1
done
