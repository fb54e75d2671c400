#!/bin/bash
# GPU-box call: the near-tie screen + 32-bit row offsets of K1's minimal mode (the build) against the commit before
# (tools/ab/libtfpose_decode_head.so): GPU suite first (near-tie counts must stay bit-identical), then K1 alone and the
# pipelined minimal-mode rate for both.  bash tools/gpu_round_check7.sh [prefix]
P=${1:-r02l}
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/${P}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${P}_pytest.log
python tools/k1_ab.py tools/ab/libtfpose_decode_head.so tf-pose-estimation_b200/libtfpose_decode.so > gpurun_out/${P}_k1_ab_head.log 2>&1
grep -E "^c[2-5] (dense_ctas128|dense_ctas148|minimal_ctas148|minimal_no_early|dense GB)" gpurun_out/${P}_k1_ab_head.log
for lib in tools/ab/libtfpose_decode_head.so tf-pose-estimation_b200/libtfpose_decode.so; do
  echo "== pipelined, $lib"; TFPOSE_DECODE_LIB=$PWD/$lib python tools/k1_sweep.py 2>&1 | grep "pipelined minimal, tail_priority 0" | tee -a gpurun_out/${P}_pipelined_minimal.log
done
python bench.py > gpurun_out/${P}_bench.json 2> gpurun_out/${P}_bench.err; echo "bench exit $?"; tail -2 gpurun_out/${P}_bench.err
python -c "
import json
l=json.load(open('gpurun_out/${P}_bench.json'))
print('value',l['value'],'ms',l['ms_per_step'],'frac',l['roofline']['frac'],'k1',l['roofline']['kernel_ms'])
print('min pipelined',l['value_minimal_mode_pipelined'],'min serial',l['value_minimal_mode'],'k1 min ms',l['roofline_minimal']['kernel_ms'])
print('e2e',l['e2e']['value'],'objects',l['e2e_objects']['value'])
for k,v in l.get('per_config',{}).items(): print(k,'value',v['value'],'ms',v['ms_per_step'],'min pipelined',v['value_minimal_mode_pipelined'],'frac',v['k1_roofline_frac'])
"
