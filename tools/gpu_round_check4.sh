#!/bin/bash
# GPU-box call: GPU suite, the checked-build runs with their negative controls, K1 A/B with the band sweep, pipelined
# sweeps (contexts in flight x tail_priority), K2 against rejection / capacity, the bench line.  bash tools/gpu_round_check4.sh [prefix]
P=${1:-r02g}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${P}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${P}_pytest.log
bash tools/checked_run.sh
python tools/k1_ab.py > gpurun_out/${P}_k1_ab.log 2>&1; grep -E "^c[2-5] (dense_ctas128|minimal|dense GB)" gpurun_out/${P}_k1_ab.log
python tools/k1_sweep.py > gpurun_out/${P}_k1_sweep.log 2>&1; grep -v "K1 dense alone" gpurun_out/${P}_k1_sweep.log
python tools/k2_prefilter_times.py > gpurun_out/${P}_k2_prefilter.log 2>&1; grep "k2_prefilter 1" gpurun_out/${P}_k2_prefilter.log | awk 'NR%2==0'
python bench.py > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${P}_bench.err; python -c "
import json
l=json.load(open('gpurun_out/${P}_bench.json'))
print('value',l['value'],'ms',l['ms_per_step'],'frac',l['roofline']['frac'],'k1',l['roofline']['kernel_ms'],'traffic',l['roofline']['traffic'])
print('min pipelined',l['value_minimal_mode_pipelined'],'min serial',l['value_minimal_mode'],'k1 min ms',l['roofline_minimal']['kernel_ms'],'tflops',l['roofline_minimal']['achieved_tflops'])
print('e2e',l['e2e']['value'],'objects',l['e2e_objects']['value'])
for k,v in l.get('per_config',{}).items(): print(k,'value',v['value'],'ms',v['ms_per_step'],'min pipelined',v['value_minimal_mode_pipelined'],'frac',v['k1_roofline_frac'],'stage',v['stage_ms'])
"
