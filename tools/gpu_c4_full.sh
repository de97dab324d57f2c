#!/bin/bash
# one --set full capture of the training GEMM (4th steady-state launch of gemm_tc_kernel after warm-up)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel --launch-skip 100 --launch-count 2 -o gpurun_out/full_c4_gemm -f python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c4_full.log 2>&1
echo "rc=$?"; tail -2 gpurun_out/ncu_c4_full.log | cut -c1-200
ncu -i gpurun_out/full_c4_gemm.ncu-rep --page raw --csv > gpurun_out/full_c4_gemm_raw.csv 2>/dev/null; ls -la gpurun_out/full_c4_gemm*
