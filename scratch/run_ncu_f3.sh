#!/bin/bash
SECS="--section SpeedOfLight --section MemoryWorkloadAnalysis --section ComputeWorkloadAnalysis --section LaunchStats --section Occupancy --section SchedulerStats --section WarpStateStats --section InstructionStats"
MET="--metrics dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_tensor.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum"
timeout 300 ncu $SECS $MET --clock-control none -k regex:fused3_kernel -s 4 -c 1 -o gpurun_out/${1}_fused3_train -f python scratch/ncu_target.py > gpurun_out/${1}_ncu_train.log 2>&1; echo "ncu train rc=$?"
timeout 300 ncu $SECS $MET --clock-control none -k regex:fused3_kernel -s 8 -c 1 -o gpurun_out/${1}_fused3_score -f python scratch/ncu_target.py --score > gpurun_out/${1}_ncu_score.log 2>&1; echo "ncu score rc=$?"
