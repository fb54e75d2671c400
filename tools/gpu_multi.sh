#!/bin/bash
# Multi-GPU call (gpurun --gpus N -- 'bash tools/gpu_multi.sh N [prefix]'): the bench under torchrun exactly as the
# driver launches it (both arms), the per-rank H2D calibration, and the GPU suite (its two-GPU test needs N >= 2).
N=${1:-2}; P=${2:-r02}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 bench.py --gpus $N > gpurun_out/${P}_bench_n$N.json 2> gpurun_out/${P}_bench_n$N.err; echo "bench N=$N exit $?"
python -c "
import json
l=json.load(open('gpurun_out/${P}_bench_n$N.json'))
print('value',l['value'],'ms',l['ms_per_step'],'e2e',l['e2e']['value'],'objects',l['e2e_objects']['value'],'min pipelined',l['value_minimal_mode_pipelined'])
print('per_rank ms',[round(r['ms_per_step'],4) for r in l['per_rank']],'enqueue',[round(r['host_enqueue_ms_per_step'],3) for r in l['per_rank']])
for k,v in l.get('per_config',{}).items(): print(k,v['scaling'],'frames/gpu',v['frames_per_gpu'],'value',v['value'],'ms',v['ms_per_step'],'per_rank',[round(r['ms_per_step'],4) for r in v['per_rank']])
print('clocks',l['clocks'])
" || tail -5 gpurun_out/${P}_bench_n$N.err
$TR --master-port 29512 bench.py --gpus $N --impl reference --steps 3 --warmup 1 > gpurun_out/${P}_bench_reference_n$N.json 2> gpurun_out/${P}_bench_reference_n$N.err; echo "reference arm N=$N exit $?"; cut -c1-200 gpurun_out/${P}_bench_reference_n$N.json
$TR --master-port 29513 tools/h2d_calib.py > gpurun_out/${P}_h2d_calib_n$N.json 2> gpurun_out/${P}_h2d_calib_n$N.err; echo "h2d calib exit $?"; cat gpurun_out/${P}_h2d_calib_n$N.json
python tools/h2d_calib.py > gpurun_out/${P}_h2d_calib_n1_on_this_box.json 2>/dev/null; cat gpurun_out/${P}_h2d_calib_n1_on_this_box.json
[ "${SKIP_PYTEST:-0}" = 1 ] || { python -m pytest tests -m gpu -q > gpurun_out/${P}_pytest_n$N.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${P}_pytest_n$N.log; }
