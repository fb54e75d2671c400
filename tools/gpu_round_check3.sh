#!/bin/bash
# GPU-box call: GPU suite, K2 early-rejection timing, ncu --set full of K1 (dense and minimal), launch list of bench.py.
# bash tools/gpu_round_check3.sh [prefix]
P=${1:-r02e}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${P}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/${P}_pytest.log
python tools/k2_prefilter_times.py 2>&1 | tee gpurun_out/${P}_k2_prefilter.log
python tools/k1_sweep.py c4 2>&1 | tee gpurun_out/${P}_k1_sweep_c4.log
# the profiled commands exit 0 without ncu first
python tools/ncu_target.py 1 3 1 2 > gpurun_out/${P}_ncu_target_plain.log 2>&1 && python tools/ncu_target.py 1 3 0 2 >> gpurun_out/${P}_ncu_target_plain.log 2>&1 || { echo "ncu target fails without ncu"; tail -5 gpurun_out/${P}_ncu_target_plain.log; }
for mode in 1 0; do
  name=$([ $mode = 1 ] && echo dense || echo minimal)
  ncu --set full --clock-control none --import-source on -k regex:k1_fused_f4 --launch-skip 3 --launch-count 1 -f \
      -o gpurun_out/${P}_k1_$name python tools/ncu_target.py 1 3 $mode 2 > gpurun_out/${P}_ncu_$name.log 2>&1
  echo "ncu $name exit $?"
  ncu -i gpurun_out/${P}_k1_$name.ncu-rep --page raw --csv > gpurun_out/${P}_k1_${name}_raw.csv 2>/dev/null
done
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-per-config > gpurun_out/${P}_bench_short.json 2> gpurun_out/${P}_bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${P}_launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-per-config > gpurun_out/${P}_ncu_bench.log 2>&1
echo "launch list exit $?"; wc -l gpurun_out/${P}_launches_bench.csv
