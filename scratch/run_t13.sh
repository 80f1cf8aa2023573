#!/bin/bash
# round-1 evidence run: GPU tests, bench line, ncu launch list of bench.py, ncu --set full of the training kernel
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x > gpurun_out/t13_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t13_pytest.log
python bench.py > gpurun_out/t13_bench.json 2> gpurun_out/t13_bench.err; echo "bench rc=$?"
python scratch/dbg9.py > gpurun_out/t13_dbg.log 2>&1
K='regex:fused2_kernel|adam_img_kernel|fused_train_kernel'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 60 --csv --log-file gpurun_out/t13_launches.csv \
  python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/t13_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:fused2_kernel -s 4 -c 1 -o gpurun_out/t13_fused2 -f \
  python scratch/ncu_target.py > gpurun_out/t13_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:adam_img_kernel -s 4 -c 1 -o gpurun_out/t13_adam -f \
  python scratch/ncu_target.py > gpurun_out/t13_ncu_adam.log 2>&1; echo "ncu adam rc=$?"
tail -3 gpurun_out/t13_pytest.log; cat gpurun_out/t13_bench.json | cut -c1-600
