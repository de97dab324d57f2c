#!/bin/bash
# ncu --set full of the warp-specialised training GEMM on the microbenchmark (after the same command ran clean)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
mkdir -p gpurun_out
GEMM_BWDW=0 GEMM_SHAPES=128x128 timeout -s KILL 120 python tools/ubench_gemm_ws.py > gpurun_out/r2_gemm_ubench.txt 2>&1; rc=$?; echo "ubench rc=$rc"; tail -8 gpurun_out/r2_gemm_ubench.txt
[ $rc -eq 0 ] || exit 1
GEMM_BWDW=0 GEMM_SHAPES=128x128 timeout -s KILL 300 ncu --set full --clock-control none --import-source on -k regex:gemm_ws_kernel -s 4 -c 1 -o gpurun_out/r2_gemm_ws -f python tools/ubench_gemm_ws.py > gpurun_out/r2_gemm_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/r2_gemm_ncu.log
