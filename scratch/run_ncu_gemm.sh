#!/bin/bash
set -x
cd /root/repo
timeout 100 python scratch/cfg5_one.py bf16x3 || exit 1
# second step's first three GEMMs: fwd layer 1 (K=400), fwd layer 2 (K=1024), fwd layer 3
timeout 600 ncu --set full --clock-control none --import-source on -k regex:umma_gemm --launch-skip 26 -c 3 -o gpurun_out/ncu_gemm_fwd -f python scratch/cfg5_one.py bf16x3 > gpurun_out/ncu_gemm.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:umma_gemm --launch-skip 26 -c 26 --csv --log-file gpurun_out/ncu_gemm_launches.csv python scratch/cfg5_one.py bf16x3 > /dev/null 2>&1
tail -3 gpurun_out/ncu_gemm.log
