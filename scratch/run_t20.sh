#!/bin/bash
# launch list of bench.py (our kernels only) + bench line, after the bench has exited 0 without ncu
python bench.py > gpurun_out/t20_bench.json 2> gpurun_out/t20_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/t20_bench_ref.json 2> gpurun_out/t20_bench_ref.err; echo "ref rc=$?"
K='regex:fused3_kernel|adam_img_kernel|hist_kernel|preprocess_kernel'
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 80 --csv --log-file gpurun_out/t20_launches.csv \
  python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/t20_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:adam_img_kernel -s 4 -c 1 -o gpurun_out/t20_adam -f \
  python scratch/ncu_target.py > gpurun_out/t20_ncu_adam.log 2>&1; echo "ncu adam rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:fused3_kernel -s 4 -c 1 -o gpurun_out/t20_fused3_full -f \
  python scratch/ncu_target.py > gpurun_out/t20_ncu_full.log 2>&1; echo "ncu fused3 full rc=$?"
cut -c1-300 gpurun_out/t20_bench.json; cut -c1-300 gpurun_out/t20_bench_ref.json
