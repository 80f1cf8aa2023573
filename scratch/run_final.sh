#!/bin/bash
# Round-1 evidence: bench (both arms) first, then -- only after they exited 0 without ncu -- the launch list and the captures.
cd /root/repo
T=${1:-t50}
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
K='regex:fused3_kernel|adam_img_kernel|hist_kernel|umma_gemm_kernel|combine_kernel'
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/${T}_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:fused3_kernel -s 4 -c 1 -o gpurun_out/${T}_fused3_full -f \
  python scratch/ncu_target.py > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu fused3 full rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:adam_img_kernel -s 4 -c 1 -o gpurun_out/${T}_adam -f \
  python scratch/ncu_target.py > gpurun_out/${T}_ncu_adam.log 2>&1; echo "ncu adam rc=$?"
SECS="--section SpeedOfLight --section MemoryWorkloadAnalysis --section ComputeWorkloadAnalysis --section LaunchStats --section Occupancy --section SchedulerStats --section WarpStateStats --section InstructionStats"
MET="--metrics dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum"
timeout 300 ncu $SECS $MET --clock-control none -k regex:fused3_kernel -s 8 -c 1 -o gpurun_out/${T}_fused3_score -f \
  python scratch/ncu_target.py --score > gpurun_out/${T}_ncu_score.log 2>&1; echo "ncu score rc=$?"
# layered GEMM at cfg-5: forward layer 2 (K=1024) and the first dW of the second step
timeout 300 ncu --set full --clock-control none --import-source on -k regex:umma_gemm --launch-skip 27 -c 1 -o gpurun_out/${T}_gemm_fwd -f \
  python scratch/cfg5_one.py bf16x3 > gpurun_out/${T}_ncu_gemm.log 2>&1; echo "ncu gemm fwd rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:umma_gemm --launch-skip 35 -c 1 -o gpurun_out/${T}_gemm_dw -f \
  python scratch/cfg5_one.py bf16x3 >> gpurun_out/${T}_ncu_gemm.log 2>&1; echo "ncu gemm dw rc=$?"
cut -c1-300 gpurun_out/${T}_bench.json; cut -c1-300 gpurun_out/${T}_bench_ref.json
