#!/bin/bash
# quick GPU iteration: x3-mode parity tests + one bench line (fields of interest only)
tag=${1:-q}
timeout 240 python -m pytest tests -m gpu -q -x -k "bf16 or config1 or smoke or mlp" > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$tag.log | cut -c1-300
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline ${@:2} > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err; echo "bench rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_$tag.json'));print('rays/s',round(d['value']),'e2e',round(d['e2e']['value']),d['kernels_ms_per_step'],'frac',round(d['roofline']['frac'],4))"; tail -3 gpurun_out/bench_$tag.err
