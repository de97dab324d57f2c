#!/bin/bash
mkdir -p gpurun_out
for rep in 1 2; do for fl in 48 0 32 16; do
  UCN_GEMM_DBG=$fl timeout 300 python bench.py --workload c4 --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/ab_c4_$fl.json 2> gpurun_out/ab_c4_$fl.err
  python -c "
import json;d=json.load(open('gpurun_out/ab_c4_$fl.json'));print('flags $fl rep $rep ms',round(d['ms_per_step'],3),{k:round(v,3) for k,v in d['kernels_ms_per_step'].items()})"
done; done
