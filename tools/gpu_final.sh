#!/bin/bash
# round-end evidence at HEAD: parity tests, smoke, c2 bench + reference arm + ncu launch list, c4 line + launch list
tag=${1:-f}
bash tools/gpu_verify.sh $tag
timeout 600 python bench.py --workload c4 --steps 30 --warmup 5 > gpurun_out/bench_c4_$tag.json 2> gpurun_out/bench_c4_$tag.err; echo "c4 rc=$?"; cut -c1-200 gpurun_out/bench_c4_$tag.json
bash tools/gpu_c4_ncu.sh > gpurun_out/c4_launch_summary_$tag.txt 2>&1; head -14 gpurun_out/c4_launch_summary_$tag.txt | cut -c1-120
