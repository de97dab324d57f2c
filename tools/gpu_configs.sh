#!/bin/bash
tag=${1:-m}
mkdir -p gpurun_out
timeout 600 python bench.py --workload c4 --steps 20 --warmup 3 > gpurun_out/bench_c4_$tag.json 2> gpurun_out/bench_c4_$tag.err; echo "c4 rc=$?"; cut -c1-1500 gpurun_out/bench_c4_$tag.json
timeout 600 python bench.py --workload c4 --impl reference --steps 2 --warmup 1 > gpurun_out/bench_c4_ref_$tag.json 2> gpurun_out/bench_c4_ref_$tag.err; echo "c4 ref rc=$?"
timeout 600 python bench.py --workload c1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c1_$tag.json 2> gpurun_out/bench_c1_$tag.err; echo "c1 rc=$?"; cut -c1-300 gpurun_out/bench_c1_$tag.json
timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_$tag.json 2> gpurun_out/bench_c5_$tag.err; echo "c5 rc=$?"; cut -c1-300 gpurun_out/bench_c5_$tag.json
