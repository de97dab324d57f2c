#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 300 python -m pytest tests/test_gpu_grad.py -m gpu -q -x -k "selection or train_rays or sampler" > gpurun_out/pytest_r2i.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_r2i.log | cut -c1-300
timeout -s KILL 300 python bench.py --workload c4 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_r2i.json 2> gpurun_out/bench_c4_r2i.err; echo "c4 rc=$?"; tail -2 gpurun_out/bench_c4_r2i.err | cut -c1-300
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c4_r2i.json'))
print(d['value'], d['ms_per_step'], d['kernels_ms_per_step'], d['gpu_launches'], d['loss_first_last'])
PY
