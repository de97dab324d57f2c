#!/bin/bash
# GPU visit for the training path: gradient parity tests (+ optional config-4 bench)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_grad.py -q -m gpu 2>&1 | tail -40 > gpurun_out/pytest_grad.log
cat gpurun_out/pytest_grad.log
if [ -n "$BENCH4" ]; then
  timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
  echo "c4 rc=$?"; tail -3 gpurun_out/bench_c4.err; cat gpurun_out/bench_c4.json
fi
