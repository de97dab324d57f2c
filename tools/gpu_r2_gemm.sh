#!/bin/bash
# training-GEMM iteration: A/B microbenchmark, the linear-primitive and gradient tests, a short config-4 bench
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout -s KILL 120 python tools/ubench_gemm_ws.py > gpurun_out/r2_gemm_ubench.txt 2>&1; echo "ubench rc=$?"; cat gpurun_out/r2_gemm_ubench.txt | tail -32
timeout -s KILL 400 python -m pytest tests/test_gpu_grad.py -m gpu -x -q > gpurun_out/r2_gemm_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_gemm_pytest.log | cut -c1-250
timeout -s KILL 200 python bench.py --workload c4 --steps 20 --warmup 3 > gpurun_out/r2_gemm_bench.json 2> gpurun_out/r2_gemm_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/r2_gemm_bench.err | cut -c1-300
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/r2_gemm_bench.json'))
    print({k:d[k] for k in ('value','ms_per_step','kernels_ms_per_step','gpu_launches')})
except Exception as e: print("no bench json", e)
PY
