#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/t14_smoke.log 2>&1; echo "smoke rc=$?"
python -m pytest tests -m gpu -q -x > gpurun_out/t14_pytest.log 2>&1; echo "pytest rc=$?"
python bench.py > gpurun_out/t14_bench.json 2> gpurun_out/t14_bench.err; echo "bench rc=$?"
K='regex:fused2_kernel|adam_img_kernel|hist_kernel'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 60 --csv --log-file gpurun_out/t14_launches.csv \
  python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/t14_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fused2_kernel -s 4 -c 1 -o gpurun_out/t14_fused2 -f \
  python scratch/ncu_target.py > gpurun_out/t14_ncu_full.log 2>&1; echo "ncu full rc=$?"
if [ ! -f gpurun_out/t14_fused2.ncu-rep ]; then
timeout 300 ncu --section SpeedOfLight --section MemoryWorkloadAnalysis --section ComputeWorkloadAnalysis --section LaunchStats --section Occupancy --section SchedulerStats --section WarpStateStats --section InstructionStats \
  --metrics dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tensor.sum \
  --clock-control none -k regex:fused2_kernel -s 4 -c 1 -o gpurun_out/t14_fused2_nosrc -f \
  python scratch/ncu_target.py > gpurun_out/t14_ncu_full2.log 2>&1; echo "ncu full (no source counters) rc=$?"
fi
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fused2_kernel -s 8 -c 1 -o gpurun_out/t14_score -f \
  python scratch/ncu_target.py --score > gpurun_out/t14_ncu_score.log 2>&1; echo "ncu score rc=$?"
tail -3 gpurun_out/t14_pytest.log; cat gpurun_out/t14_bench.json | cut -c1-300
