#!/bin/bash
# One GPU-box call of the development loop: GPU test-suite, K1 A/B timing against tools/ab/libtfpose_decode_base.so,
# the bench line (c2 + per_config) and the reference arm.  Run from the repo root: bash tools/gpu_round_check.sh
# (outputs under gpurun_out/, prefix ${P}_; pass another prefix as $1)
P=${1:-r02c}
python -m pytest tests -m gpu -q -s > gpurun_out/${P}_pytest.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/${P}_pytest.log
python tools/k1_ab.py > gpurun_out/${P}_k1_ab.log 2>&1; grep -v "^JSON" gpurun_out/${P}_k1_ab.log
python bench.py > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${P}_bench.err; python -c "
import json
l=json.load(open('gpurun_out/${P}_bench.json'))
print('value',l['value'],'ms',l['ms_per_step'],'frac',l['roofline']['frac'],'k1',l['roofline']['kernel_ms'])
print('min pipelined',l['value_minimal_mode_pipelined'],'min serial',l['value_minimal_mode'],'roofline_minimal',l['roofline_minimal'])
print('e2e',l['e2e']['value'],'objects',l['e2e_objects'],'dense',l.get('e2e_dense_maps'))
print('per_rank',l['per_rank']); print('clocks',l['clocks'])
for k,v in l.get('per_config',{}).items(): print(k,'value',v['value'],'ms',v['ms_per_step'],'serial',v['value_one_call_at_a_time'],'frac',v['k1_roofline_frac'],'stage',v['stage_ms'],'peaks/frame',v['peaks_per_frame'])
c=l.get('cpu_baseline'); print('cpu',c and {k:c[k] for k in c if k not in ('parts',)})
"
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${P}_bench_reference.json 2> gpurun_out/${P}_bench_reference.err; echo "ref exit $?"; cut -c1-600 gpurun_out/${P}_bench_reference.json
