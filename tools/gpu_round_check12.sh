#!/bin/bash
# GPU-box call: staged strip resize: parity tests, times for strip heights, ncu --set full of the staged kernel.
P=${1:-r03f}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_preproc.py -m gpu -x -q 2>&1 | tail -4
for ty in 0 16; do echo "== TY $ty (0 = the launcher's model)"; TFPD_RESIZE_TY=$ty timeout 300 python tools/preproc_times.py 2>&1 | grep -E "staged" | tee -a gpurun_out/${P}_preproc_times.log; done
timeout 300 python tools/preproc_times.py 2>&1 | grep -E "resize_cubic (strip|tiled|per)" | tee -a gpurun_out/${P}_preproc_times.log
timeout 300 python tools/preproc_times.py 64 720 1280 2>&1 | grep -E "resize_cubic (staged|strip)" | tee -a gpurun_out/${P}_preproc_times.log
timeout 300 python tools/preproc_times.py 16 2>&1 | grep -E "resize_cubic (staged|strip)" | tee -a gpurun_out/${P}_preproc_times.log
python tools/ncu_preproc.py 0 > /dev/null 2>&1 || { echo "target fails"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_resize_cubic_u8 --launch-skip 1 --launch-count 1 -f \
    -o gpurun_out/${P}_resize_staged python tools/ncu_preproc.py 0 > gpurun_out/${P}_ncu_staged.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/${P}_resize_staged.ncu-rep --page raw --csv > gpurun_out/${P}_resize_staged_raw.csv 2>/dev/null
python tools/ncu_pick.py gpurun_out/${P}_resize_staged_raw.csv --csv gpurun_out/${P}_resize_ncu_full_summary.csv | grep -E "time_duration|inst_executed.sum|issue_active|warps_active|pipe_|wavefronts_mem_shared|bank_conflicts|grid_size|block_size|shared_mem|registers|stall_(barrier|long|short|wait|math|not_sel|mio)"
