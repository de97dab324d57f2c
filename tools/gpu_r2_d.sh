#!/bin/bash
# K3 iteration: parity tests that run the tensor-core kernel + a short bench + CTA-0 timeline (hard timeouts: a deadlocked
# kernel must not burn the lease)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 60 python tools/trace_x3.py 0 420 > gpurun_out/r2d_trace.txt 2>&1; echo "trace rc=$?"; head -1 gpurun_out/r2d_trace.txt
timeout -s KILL 150 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bf16 or config or host_sweep or row_band or operand or round1" > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2d_pytest.log
tail -12 gpurun_out/r2d_pytest.log | cut -c1-250
timeout -s KILL 120 python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/r2d_bench.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2d_bench.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['roofline']['frac'], d['kernels_ms_per_step'], d['parity'])
PY
