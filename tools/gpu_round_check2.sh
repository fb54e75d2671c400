#!/bin/bash
# GPU-box call after the minimal-mode queue and the two-per-SM K3: GPU suite, K1 A/B (with the band sweep), CTA-count and
# pipeline-depth sweeps, K3 one-vs-two CTAs per SM on crowded frames, the bench line.  bash tools/gpu_round_check2.sh [prefix]
P=${1:-r02d}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s > gpurun_out/${P}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/${P}_pytest.log
python tools/k1_ab.py > gpurun_out/${P}_k1_ab.log 2>&1; grep -v "^JSON" gpurun_out/${P}_k1_ab.log
python tools/k1_sweep.py > gpurun_out/${P}_k1_sweep.log 2>&1; cat gpurun_out/${P}_k1_sweep.log
TFPD_K3_MINB=1 python tools/k1_sweep.py c4 > gpurun_out/${P}_k1_sweep_k3_one_per_sm.log 2>&1; cat gpurun_out/${P}_k1_sweep_k3_one_per_sm.log
python bench.py > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${P}_bench.err; python -c "
import json
l=json.load(open('gpurun_out/${P}_bench.json'))
print('value',l['value'],'ms',l['ms_per_step'],'frac',l['roofline']['frac'],'k1',l['roofline']['kernel_ms'])
print('min pipelined',l['value_minimal_mode_pipelined'],'min serial',l['value_minimal_mode'],'k1 min ms',l['roofline_minimal']['kernel_ms'],'tflops',l['roofline_minimal']['achieved_tflops'])
print('e2e',l['e2e']['value'],'objects',l['e2e_objects']['value'])
for k,v in l.get('per_config',{}).items(): print(k,'value',v['value'],'ms',v['ms_per_step'],'min pipelined',v['value_minimal_mode_pipelined'],'frac',v['k1_roofline_frac'],'stage',v['stage_ms'])
"
