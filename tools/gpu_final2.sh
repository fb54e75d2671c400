#!/bin/bash
# Closing records of the round's last code on one GPU: tools/gpu_final.sh (GPU suite, checked build with its negative
# controls, both bench arms) + the host side under ASan/UBSan with a live GPU + the pre-processing times.
P=${1:-r03_final}
bash tools/gpu_final.sh $P
PRE="$(gcc -print-file-name=libasan.so):$(gcc -print-file-name=libubsan.so)"
mkdir -p gpurun_out/r02_sanitizers
LD_PRELOAD=$PRE TFPOSE_DECODE_LIB=$PWD/tools/ab/libtfpose_decode_asan.so ASAN_OPTIONS=protect_shadow_gap=0:detect_leaks=0:halt_on_error=0:log_path=$PWD/gpurun_out/r02_sanitizers/host_asan_report \
  UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=0:log_path=$PWD/gpurun_out/r02_sanitizers/host_ubsan_report \
  timeout 900 python -m pytest tests/test_gpu_seams.py tests/test_gpu_parity.py tests/test_preproc.py tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r02_sanitizers/host_asan_pytest.log 2>&1
echo "host ASan/UBSan with a GPU: exit $? | $(tail -1 gpurun_out/r02_sanitizers/host_asan_pytest.log) | reports: $(ls gpurun_out/r02_sanitizers | grep -cE 'host_(asan|ubsan)_report')"
timeout 300 python tools/preproc_times.py > gpurun_out/${P}_preproc_times.log 2>&1; grep -E "resize_cubic|quantize" gpurun_out/${P}_preproc_times.log
