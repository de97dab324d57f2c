#!/bin/bash
# A/B of the per-frame barrier placement at N GPUs (usage: gpu_r2_bar.sh N)
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
for b in inline side inline side; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 3 --no-extras --no-cpu-baseline --barrier $b > gpurun_out/bar_${b}_${N}.json 2> gpurun_out/bar_${b}_${N}.err; echo "$b rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/bar_${b}_${N}.json').read().strip().splitlines()[-1])
print('$b', {k:d.get(k) for k in ("value","ms_per_step","host_submit_ms_per_step","gather_ms","per_rank_kernel_ms","checks")}, (d.get("e2e") or {}).get("value"))
PY
done
