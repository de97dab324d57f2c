#!/bin/bash
# gather-focused GPU visit: parity tests for K2, bench line, optional ncu --set full of the gather kernel at C2
tag=${1:-g}
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -q -x -k "gather or whole_frame or config1 or drop_in" > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_$tag.log | cut -c1-300
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_$tag.json'));print('rays/s',round(d['value']),d['kernels_ms_per_step'],d['hbm_kernels'])"
if [ -n "$NCU" ]; then timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gather' -c 1 -o gpurun_out/gfull_$tag -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_gfull_$tag.log 2>&1; echo "ncu rc=$?"; fi
