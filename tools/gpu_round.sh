#!/bin/bash
# one GPU-box visit: parity tests, bench, ncu launch list, ncu --set full of the two dominant kernels
tag=${1:-a}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/smi_$tag.txt
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$tag.log | cut -c1-300
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"; cat gpurun_out/bench_$tag.json | cut -c1-3000
timeout 300 python tools/trace_x3.py 0 400 > gpurun_out/trace_$tag.txt 2>&1; echo "trace rc=$?"; head -3 gpurun_out/trace_$tag.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch_$tag.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'mlp_x3_kernel|gather' -c 4 -o gpurun_out/full_$tag -f python bench.py --workload c1 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full_$tag.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
