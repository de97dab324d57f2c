#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras > gpurun_out/plain_r2g.log 2>&1 && \
timeout -s KILL 700 ncu --set full --clock-control none --import-source on -k regex:'mlp_x3p_kernel|gather_kernel|composite_kernel|raygen_sample_kernel|ray_term_kernel' -c 5 -o gpurun_out/full_r2a -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras > gpurun_out/ncu_full_r2a.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/ncu_full_r2a.log | cut -c1-200
