#!/bin/bash
# round 2: sharded bench at N GPUs (N = first arg, default 2) + the band test
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "row_band or wrong_device or frame_cache or default_seed" > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
tail -3 gpurun_out/r2b_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2b_bench_${N}gpu.json 2> gpurun_out/r2b_bench_${N}gpu.err; echo "bench rc=$?" >> gpurun_out/r2b_bench_${N}gpu.err
tail -c 1500 gpurun_out/r2b_bench_${N}gpu.err; head -c 3000 gpurun_out/r2b_bench_${N}gpu.json
