#!/bin/bash
# Evidence after the scoring-pass rework: bench (both arms) first, then -- only after they exited 0 without ncu -- the launch list
# and the captures of the two fused3 instantiations (training step, scoring pass).
cd /root/repo
T=${1:-t60}
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
K='regex:fused3_kernel|adam_img_kernel|hist_kernel|umma_gemm_kernel|combine_kernel'
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/${T}_launches.csv \
  python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/${T}_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fused3_kernel -s 8 -c 1 -o gpurun_out/${T}_fused3_score -f \
  python scratch/ncu_target.py --score > gpurun_out/${T}_ncu_score.log 2>&1; echo "ncu score full rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:fused3_kernel -s 4 -c 1 -o gpurun_out/${T}_fused3_full -f \
  python scratch/ncu_target.py > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu fused3 train full rc=$?"
cut -c1-300 gpurun_out/${T}_bench.json; cut -c1-300 gpurun_out/${T}_bench_ref.json
