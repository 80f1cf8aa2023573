#!/bin/bash
# ncu captures of round 2 (one GPU).  $1 = tag.  The plain run comes first; each capture only after it exited 0.
T=${1:-r02}
python bench/ncu_target.py > gpurun_out/${T}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${T}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${T}_launches_target.csv python bench/ncu_target.py > gpurun_out/${T}_ncu_launch.log 2>&1; echo "launch list rc=$?"
for K in "fused3_kernel" "adam_img_kernel" "hist_kernel" "minmax_kernel" "loss_map_kernel" "preprocess_kernel" "ingest_f16_kernel" "rs_scatter_kernel" "seq_cumsum2_kernel|dscan_apply_kernel"; do
  N=$(echo "$K" | tr -c 'a-zA-Z0-9_' '_')
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:$K" -c 4 -f -o gpurun_out/${T}_${N} python bench/ncu_target.py > gpurun_out/${T}_ncu_${N}.log 2>&1; echo "ncu $K rc=$?"
  # gpurun_out/ may carry 64 MiB back: keep the text pages, drop the report (the training / scoring kernel's report is kept)
  ncu -i gpurun_out/${T}_${N}.ncu-rep --page details > gpurun_out/${T}_${N}_details.txt 2>/dev/null
  python bench/ncu_summary.py gpurun_out/${T}_${N}.ncu-rep > gpurun_out/${T}_${N}_raw_summary.txt 2>/dev/null
  [ "$K" = "fused3_kernel" ] || rm -f gpurun_out/${T}_${N}.ncu-rep
done
