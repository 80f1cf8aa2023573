#!/bin/bash
cd /root/repo
timeout 100 python scratch/cfg5_one.py bf16x3 || exit 1
# second step: launch 27 = fwd layer 2 (K=1024), 35 = first dW
timeout 600 ncu --set full --clock-control none --import-source on -k regex:umma_gemm --launch-skip 27 -c 1 -o gpurun_out/ncu_gemm_fwd2 -f python scratch/cfg5_one.py bf16x3 > gpurun_out/ncu_gemm.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:umma_gemm --launch-skip 35 -c 1 -o gpurun_out/ncu_gemm_dw -f python scratch/cfg5_one.py bf16x3 >> gpurun_out/ncu_gemm.log 2>&1
tail -3 gpurun_out/ncu_gemm.log
