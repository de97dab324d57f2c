#!/bin/bash
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
CUDA_LAUNCH_BLOCKING=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 2 --warmup 3 --no-extras > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?" >> gpurun_out/r2c_bench.err
grep -v "^frame\|^$" gpurun_out/r2c_bench.err | head -40 | cut -c1-300; head -c 1000 gpurun_out/r2c_bench.json
