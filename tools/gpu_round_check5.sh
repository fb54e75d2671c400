#!/bin/bash
# GPU-box call: dense K1 with / without the rotated PAF staging (tools/ab/libtfpose_decode_prev.so = HEAD without it),
# ncu --set full of K1 dense and minimal (final kernels), K2 against max_cands, GPU suite.  bash tools/gpu_round_check5.sh [prefix]
P=${1:-r02i}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${P}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${P}_pytest.log
python tools/k1_ab.py tools/ab/libtfpose_decode_prev.so tf-pose-estimation_b200/libtfpose_decode.so > gpurun_out/${P}_k1_ab_rot.log 2>&1; grep -E "^c[2-5] (dense|minimal_ctas148)" gpurun_out/${P}_k1_ab_rot.log
python tools/k1_ab.py tools/ab/libtfpose_decode_prev.so tf-pose-estimation_b200/libtfpose_decode.so > gpurun_out/${P}_k1_ab_rot2.log 2>&1; grep -E "^c[2-5] dense" gpurun_out/${P}_k1_ab_rot2.log
python tools/k2_prefilter_times.py > gpurun_out/${P}_k2_caps.log 2>&1; grep "k2_prefilter 1" gpurun_out/${P}_k2_caps.log | awk 'NR%2==0'
python tools/ncu_target.py 1 3 1 2 > gpurun_out/${P}_ncu_target_plain.log 2>&1 && python tools/ncu_target.py 1 3 0 2 >> gpurun_out/${P}_ncu_target_plain.log 2>&1 || { echo "ncu target fails without ncu"; tail -5 gpurun_out/${P}_ncu_target_plain.log; }
for mode in 1 0; do
  name=$([ $mode = 1 ] && echo dense || echo minimal)
  ncu --set full --clock-control none --import-source on -k regex:k1_fused_f4 --launch-skip 3 --launch-count 1 -f \
      -o gpurun_out/${P}_k1_$name python tools/ncu_target.py 1 3 $mode 2 > gpurun_out/${P}_ncu_$name.log 2>&1
  echo "ncu $name exit $?"
  ncu -i gpurun_out/${P}_k1_$name.ncu-rep --page raw --csv > gpurun_out/${P}_k1_${name}_raw.csv 2>/dev/null
done
python tools/ncu_pick.py gpurun_out/${P}_k1_dense_raw.csv gpurun_out/${P}_k1_minimal_raw.csv --csv gpurun_out/${P}_k1_ncu_full_summary.csv | grep -E "time_duration|bank_conflicts|wavefronts_mem_shared|inst_executed.sum|pipe_fma|issue_active|l1tex__throughput|cycles_active|cycles_elapsed|dram__bytes"
