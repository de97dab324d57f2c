#!/bin/bash
# same-box A/B of two builds of libucnerf.so on the K3 microbenchmark.  Prepare here first (the .so files are git-ignored but
# travel with gpurun):  build the new tree, cp uc-nerf_b200/libucnerf.so tools/ab/libucnerf_new.so;  git stash, build,
# cp ... tools/ab/libucnerf_old.so;  git stash pop, rebuild.  Read the MIN column: the median moves with the clock ramp.
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for v in old new old new; do
  cp tools/ab/libucnerf_$v.so uc-nerf_b200/libucnerf.so
  echo -n "$v: "; X3_SCALE=40 X3_AB="0" timeout -s KILL 200 python tools/trace_x3.py 2>&1 | tail -1
done
