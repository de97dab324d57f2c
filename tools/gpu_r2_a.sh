#!/bin/bash
# round 2, first GPU pass: parity tests (incl. the new config-2 / config-5 / band tests) + one bench line with extras
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r2a_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?" >> gpurun_out/r2a_bench.err
tail -3 gpurun_out/r2a_pytest.log; tail -c 600 gpurun_out/r2a_bench.err; head -c 1500 gpurun_out/r2a_bench.json
