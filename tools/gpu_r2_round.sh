#!/bin/bash
# one GPU-box visit (round 2): parity tests, smoke, bench (+ reference arm), ncu launch list (the --set full capture is tools/gpu_r2_g.sh: one ncu per visit)
tag=${1:-a}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_r2$tag.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_r2$tag.log | cut -c1-300
timeout -s KILL 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_r2$tag.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke_r2$tag.log | cut -c1-400
timeout -s KILL 400 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r2$tag.json 2> gpurun_out/bench_r2$tag.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench_r2$tag.json; tail -3 gpurun_out/bench_r2$tag.err | cut -c1-300
timeout -s KILL 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_r2$tag.json 2> gpurun_out/bench_ref_r2$tag.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/bench_ref_r2$tag.json
timeout -s KILL 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > gpurun_out/plain_r2$tag.log 2>&1 && \
timeout -s KILL 500 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r2$tag.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > gpurun_out/ncu_launch_r2$tag.log 2>&1; echo "ncu launches rc=$?"
ls -la gpurun_out | tail -8
