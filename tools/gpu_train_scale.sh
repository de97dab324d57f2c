#!/bin/bash
# N-GPU visit for config 4: training step with the NCCL gradient all-reduce
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --workload c4 --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_c4_${N}gpu.json 2> gpurun_out/bench_c4_${N}gpu.err; echo "c4 rc=$?"
tail -3 gpurun_out/bench_c4_${N}gpu.err; grep '"metric"' gpurun_out/bench_c4_${N}gpu.json | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:d[k] for k in ['value','n_gpus','ms_per_step','kernels_ms_per_step','loss_first_last']})"
