#!/bin/bash
# one GPU-box visit: parity tests, smoke, bench (+ reference arm), ncu launch list, ncu --set full of every hot-path kernel at C2
tag=${1:-a}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/smi_$tag.txt
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$tag.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_$tag.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke_$tag.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"; cut -c1-2500 gpurun_out/bench_$tag.json; tail -3 gpurun_out/bench_$tag.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref_$tag.err; echo "ref rc=$?"; cut -c1-900 gpurun_out/bench_ref_$tag.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$tag.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch_$tag.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'mlp_x3_kernel|gather_kernel|composite_kernel|raygen_sample_kernel' -c 4 -o gpurun_out/full_$tag -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full_$tag.log 2>&1; echo "ncu full rc=$?"
# config 4 (training step), config 1 and config 5 lines; launch list of the training step
timeout 600 python bench.py --workload c4 --steps 20 --warmup 3 > gpurun_out/bench_c4_$tag.json 2> gpurun_out/bench_c4_$tag.err; echo "c4 rc=$?"; cut -c1-300 gpurun_out/bench_c4_$tag.json
timeout 600 python bench.py --workload c4 --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c4_ref_$tag.json 2> gpurun_out/bench_c4_ref_$tag.err; echo "c4 ref rc=$?"
timeout 600 python bench.py --workload c1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1_$tag.json 2> gpurun_out/bench_c1_$tag.err; echo "c1 rc=$?"
timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_$tag.json 2> gpurun_out/bench_c5_$tag.err; echo "c5 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_c4_$tag.csv python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch_c4_$tag.log 2>&1; echo "ncu c4 launches rc=$?"
ls -la gpurun_out | tail -12
