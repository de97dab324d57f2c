#!/bin/bash
# ncu launch list (durations only) of the config-4 training step, after the same command ran clean
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 200 python bench.py --workload c4 --steps 2 --warmup 3 > gpurun_out/r2_c4_plain.json 2> gpurun_out/r2_c4_plain.err; rc=$?; echo "plain rc=$rc"
[ $rc -eq 0 ] || { tail -3 gpurun_out/r2_c4_plain.err; exit 1; }
timeout -s KILL 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_c4_launches.csv python bench.py --workload c4 --steps 2 --warmup 3 > gpurun_out/r2_c4_ncu.log 2>&1; echo "ncu rc=$?"
