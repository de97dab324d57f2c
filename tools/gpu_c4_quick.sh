#!/bin/bash
# config-4 iteration: gradient parity tests + one training-step bench line
tag=${1:-c4q}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_grad.py tests/test_losses.py -m gpu -q -x > gpurun_out/pytest_$tag.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$tag.log | cut -c1-300
timeout 600 python bench.py --workload c4 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c4_$tag.json 2> gpurun_out/bench_c4_$tag.err; echo "c4 rc=$?"
python -c "
import json;d=json.load(open('gpurun_out/bench_c4_$tag.json'));print('rays/s',round(d['value']),'ms',round(d['ms_per_step'],3),d['kernels_ms_per_step'],'roofline',d['roofline'].get('achieved'),d['roofline'].get('frac'),d['roofline'].get('ms_per_launch'),d['clocks'])"; tail -3 gpurun_out/bench_c4_$tag.err
