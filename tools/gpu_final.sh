#!/bin/bash
# Records of the round's final code on one GPU: GPU suite, checked-build suite with its negative controls, the bench line
# with the CPU baseline, the reference arm.  bash tools/gpu_final.sh [prefix]
P=${1:-r02_final}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${P}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${P}_pytest.log
bash tools/checked_run.sh
python bench.py --impl reference > gpurun_out/${P}_bench_reference.json 2> gpurun_out/${P}_bench_reference.err; echo "reference arm exit $?"; cut -c1-160 gpurun_out/${P}_bench_reference.json
python bench.py > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo "bench exit $?"; tail -2 gpurun_out/${P}_bench.err
python -c "
import json
l=json.load(open('gpurun_out/${P}_bench.json'))
print('value',l['value'],'ms',l['ms_per_step'],'frac',l['roofline']['frac'],'k1',l['roofline']['kernel_ms'],'launches',l['gpu_launches'])
print('min pipelined',l['value_minimal_mode_pipelined'],'roofline_minimal',l['roofline_minimal']['achieved_tflops'],l['roofline_minimal']['pipe_fma_pct'])
print('e2e',l['e2e']['value'],'objects',l['e2e_objects']['value'],'dense',l['e2e_dense_maps']['value'])
for k,v in l['per_config'].items(): print(k,'value',v['value'],'ms',v['ms_per_step'],'frac',v['k1_roofline_frac'])
c=l['cpu_baseline']; print('cpu',c['value'],c['cores'],c['sample']); print({k:c[k] for k in c if k.startswith('stage_')})
print('clocks',l['clocks'])
"
