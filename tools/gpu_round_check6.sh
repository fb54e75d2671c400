#!/bin/bash
# GPU-box call: minimal-mode K1 with 6 warps x 4 CTAs per SM (the build) against 3 warps x 8 (tools/ab/libtfpose_decode_w3.so,
# -DTFPD_MIN_WARPS=3), then the closing validation of gpu_round_check4.sh and the host-side ASan run with a live GPU.
P=${1:-r02k}
mkdir -p gpurun_out
python tools/k1_ab.py tf-pose-estimation_b200/libtfpose_decode.so tools/ab/libtfpose_decode_w3.so > gpurun_out/${P}_k1_ab_w3.log 2>&1; grep -E "^c[2-5] minimal" gpurun_out/${P}_k1_ab_w3.log
for lib in tf-pose-estimation_b200/libtfpose_decode.so tools/ab/libtfpose_decode_w3.so; do
  echo "== pipelined, $lib"; TFPOSE_DECODE_LIB=$PWD/$lib python tools/k1_sweep.py 2>&1 | grep "pipelined minimal, tail_priority 0" | tee -a gpurun_out/${P}_pipelined_minimal_w3.log
done
bash tools/gpu_round_check4.sh $P
PRE="$(gcc -print-file-name=libasan.so):$(gcc -print-file-name=libubsan.so)"
LD_PRELOAD=$PRE TFPOSE_DECODE_LIB=$PWD/tools/ab/libtfpose_decode_asan.so ASAN_OPTIONS=protect_shadow_gap=0:detect_leaks=0:halt_on_error=0:log_path=$PWD/gpurun_out/r02_sanitizers/host_asan_report \
  UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=0:log_path=$PWD/gpurun_out/r02_sanitizers/host_ubsan_report \
  timeout 900 python -m pytest tests/test_gpu_seams.py tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider > gpurun_out/r02_sanitizers/host_asan_pytest.log 2>&1
echo "host ASan/UBSan with a GPU: exit $? | $(tail -1 gpurun_out/r02_sanitizers/host_asan_pytest.log) | reports: $(ls gpurun_out/r02_sanitizers | grep -cE 'host_(asan|ubsan)_report')"
