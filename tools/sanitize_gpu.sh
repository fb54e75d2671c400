#!/bin/bash
# GPU half of the sanitizer runs, on a B200 box:
#   compute-sanitizer --tool memcheck / racecheck / synccheck / initcheck over tools/sanitize_target.py (every
#   kernel path of the library, results checked against the oracle), once more with K3's register rows off
#   (shared-memory assembly path), and the host side of the library under ASan/UBSan with a live GPU
#   (tools/ab/libtfpose_decode_asan.so, built by tools/sanitize_cpu.sh; ASan needs protect_shadow_gap=0 next to CUDA).
# Output: gpurun_out/r02_sanitizers/*.log (copy the summaries to profiles/r02_sanitizers/).
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/gpurun_out/r02_sanitizers
mkdir -p "$OUT"
cd "$ROOT"
python tools/sanitize_target.py > "$OUT/plain_run.log" 2>&1 || { echo "target fails WITHOUT a sanitizer:"; tail -5 "$OUT/plain_run.log"; exit 1; }
tail -1 "$OUT/plain_run.log"
for tool in memcheck racecheck synccheck initcheck; do
  extra=""
  [ $tool = memcheck ] && extra="--leak-check no"
  [ $tool = racecheck ] && extra="--racecheck-report all"
  timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool $tool $extra --print-limit 30 --log-file "$OUT/gpu_$tool.log" \
      python tools/sanitize_target.py > "$OUT/gpu_${tool}_stdout.log" 2>&1
  echo "$tool: exit $? | $(tail -1 "$OUT/gpu_${tool}_stdout.log") | $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' "$OUT/gpu_$tool.log" | tail -1)"
done
TFPD_K3_REGROWS=0 TFPD_K3_PARALLEL=0 timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool memcheck --leak-check no --print-limit 30 \
    --log-file "$OUT/gpu_memcheck_k3_smem_rows.log" python tools/sanitize_target.py > "$OUT/gpu_memcheck_k3_smem_rows_stdout.log" 2>&1
echo "memcheck (K3 shared-memory rows): exit $? | $(grep 'ERROR SUMMARY' "$OUT/gpu_memcheck_k3_smem_rows.log" | tail -1)"
if [ -f tools/ab/libtfpose_decode_asan.so ]; then
  PRE="$(gcc -print-file-name=libasan.so):$(gcc -print-file-name=libubsan.so)"
  LD_PRELOAD=$PRE TFPOSE_DECODE_LIB=$ROOT/tools/ab/libtfpose_decode_asan.so \
    ASAN_OPTIONS=protect_shadow_gap=0:detect_leaks=0:halt_on_error=0:log_path=$OUT/host_asan_report \
    UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=0:log_path=$OUT/host_ubsan_report \
    timeout 900 python -m pytest tests/test_gpu_seams.py tests/test_gpu_parity.py -m gpu -q -x -p no:cacheprovider > "$OUT/host_asan_pytest.log" 2>&1
  echo "host ASan/UBSan with a GPU: exit $? | $(tail -1 "$OUT/host_asan_pytest.log") | reports: $(ls "$OUT" | grep -cE 'host_(asan|ubsan)_report')"
fi
