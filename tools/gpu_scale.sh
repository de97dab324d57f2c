#!/bin/bash
# N-GPU visit: weak-scaling bench line + strong-scaling sharded-render check
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_${N}gpu.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 tools/sharded_check.py > gpurun_out/sharded_${N}gpu.json 2> gpurun_out/sharded_${N}gpu.err; echo "sharded rc=$?"; cat gpurun_out/sharded_${N}gpu.json
