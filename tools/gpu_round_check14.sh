#!/bin/bash
# GPU-box call: ncu --set full + source page of K2 (k_limb_score) and K3 (k_assemble) on crowded frames (config 4).
P=${1:-r03j}
mkdir -p gpurun_out
python tools/ncu_target.py 1 3 1 4 > gpurun_out/${P}_ncu_target_plain.log 2>&1 || { echo "target fails"; tail -5 gpurun_out/${P}_ncu_target_plain.log; exit 1; }
for k in k_limb_score k_assemble; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 3 --launch-count 1 -f \
      -o gpurun_out/${P}_$k python tools/ncu_target.py 1 3 1 4 > gpurun_out/${P}_ncu_$k.log 2>&1
  echo "ncu $k exit $?"
  ncu -i gpurun_out/${P}_$k.ncu-rep --page raw --csv > gpurun_out/${P}_${k}_raw.csv 2>/dev/null
done
python tools/ncu_pick.py gpurun_out/${P}_k_limb_score_raw.csv gpurun_out/${P}_k_assemble_raw.csv --csv gpurun_out/${P}_k2_k3_crowded_ncu_full_summary.csv | grep -E "time_duration|inst_executed.sum|issue_active|warps_active|pipe_|grid_size|block_size|registers|sectors_pipe_lsu_mem_global_op_ld|requests_pipe_lsu_mem_global_op_ld|hit_rate|stall_"
