#!/bin/bash
# quick GPU iteration: full parity tests + one bench line (fields of interest only) [+ optional x3 trace]
tag=${1:-q}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_$tag.log | cut -c1-400
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline ${@:2} > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_$tag.json'));print('rays/s',round(d['value']),'e2e',round(d['e2e']['value']),d['kernels_ms_per_step'],'frac',round(d['roofline']['frac'],4),d['hbm_kernels'],d['clocks'])"; tail -3 gpurun_out/bench_$tag.err
if [ -n "$TRACE" ]; then timeout 200 python tools/trace_x3.py 0 500 > gpurun_out/trace_$tag.txt 2>&1; echo "trace rc=$?"; head -1 gpurun_out/trace_$tag.txt; fi
