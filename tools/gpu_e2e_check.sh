#!/bin/bash
# Is the end-to-end figure a property of the box or of the library?  Same box: bare H2D calibration, then the bench's
# c2 line (no per_config, no CPU leg) with the round-1 library and with the current one, twice each; then the launch
# lists of the final code (single-purpose target and bench itself).
P=${1:-r02o}
mkdir -p gpurun_out
python tools/h2d_calib.py | tee gpurun_out/${P}_h2d_calib.json | cut -c1-400
for rep in 1 2; do
  for lib in tools/ab/libtfpose_decode_base.so tf-pose-estimation_b200/libtfpose_decode.so; do
    TFPOSE_DECODE_LIB=$PWD/$lib python bench.py --no-per-config --no-cpu-baseline > gpurun_out/${P}_tmp.json 2> gpurun_out/${P}_tmp.err
    python -c "
import json,sys
l=json.load(open('gpurun_out/${P}_tmp.json'))
print('$lib rep $rep: value %.0f  e2e %.0f  e2e_objects %.0f  e2e_dense %.0f  min_pipelined %.0f' % (l['value'], l['e2e']['value'], l['e2e_objects']['value'], l['e2e_dense_maps']['value'], l['value_minimal_mode_pipelined']))
" | tee -a gpurun_out/${P}_e2e_ab.log || tail -3 gpurun_out/${P}_tmp.err
  done
done
python tools/e2e_breakdown.py 2>&1 | tail -12 | tee gpurun_out/${P}_e2e_breakdown.log
python tools/ncu_target.py 6 3 1 2 > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/${P}_launches_ncu_target.csv python tools/ncu_target.py 6 3 1 2 > /dev/null 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-per-config > gpurun_out/${P}_bench_short.json 2>/dev/null && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${P}_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-per-config > /dev/null 2>&1
python tools/launch_shares.py gpurun_out/${P}_launches_ncu_target.csv gpurun_out/${P}_bench_short.json | tee gpurun_out/${P}_launch_shares.txt
