#!/bin/bash
# CPU half of the sanitizer runs: AddressSanitizer + UndefinedBehaviorSanitizer builds of
#   - the oracle's C / C++ restatements (oracle/maps_oracle.c, oracle/paf_oracle.cpp),
#   - the reference's OWN tf_pose/pafprocess/pafprocess.cpp (both variants, compiled where it lies),
#   - the host side of the product library (api.cu through nvcc -Xcompiler; argument / shape / state paths),
# exercised by the CPU test-suite (tests/test_oracle.py, tests/test_host.py) with the sanitizer runtimes preloaded
# into Python.  Also records what GCC says at COMPILE time about pafprocess.cpp:152 (`int membership;` read
# uninitialised): UBSan has no check for uninitialised reads (that is MemorySanitizer, clang-only; neither clang
# nor valgrind is in this image), so the compiler diagnostic is the evidence for that line.
# Output: profiles/r02_sanitizers/cpu_*.log.  Needs /root/reference (build container).
set -u
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/profiles/r02_sanitizers
SAN=$ROOT/oracle/_san
REF=${REF:-/root/reference}/tf_pose/pafprocess
mkdir -p "$OUT" "$SAN/_ref"
SF="-O1 -g -fno-omit-frame-pointer -fsanitize=address,undefined -ffp-contract=off -fPIC -shared"
set -x
gcc $SF -std=gnu11 -o "$SAN/liboracle_maps.so" "$ROOT/oracle/maps_oracle.c" -lm || exit 1
g++ $SF -std=c++14 -o "$SAN/liboracle_paf.so" "$ROOT/oracle/paf_oracle.cpp" || exit 1
g++ $SF -Wl,-Bsymbolic -I"$REF" -o "$SAN/_ref/libpafprocess_asbuilt.so" "$REF/pafprocess.cpp" || exit 1
sed 's/int membership;/int membership = 0;/' "$REF/pafprocess.cpp" | g++ $SF -Wl,-Bsymbolic -I"$REF" -x c++ -o "$SAN/_ref/libpafprocess_upstream.so" - || exit 1
# compile-time diagnostic for pafprocess.cpp:152
g++ -O2 -Wall -Wextra -Wuninitialized -Wmaybe-uninitialized -I"$REF" -c "$REF/pafprocess.cpp" -o /dev/null 2> "$OUT/cpu_gcc_warnings_pafprocess.log"
grep -n "membership" "$OUT/cpu_gcc_warnings_pafprocess.log" | head -5
# host side of the product library with ASan/UBSan (device code unchanged)
( cd "$ROOT/tf-pose-estimation_b200/csrc" && nvcc -gencode arch=compute_100a,code=sm_100a -O1 -g -std=c++17 -lineinfo \
    -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -Xcompiler -fsanitize=address -Xcompiler -fsanitize=undefined -Xcompiler -fno-omit-frame-pointer \
    -I../../include -shared -o "$ROOT/tools/ab/libtfpose_decode_asan.so" \
    api.cu k1_fused.cu k_generic.cu k_peaks.cu k2_limbs.cu k3_assemble.cu k_draw.cu k_preproc.cu k_noalloc.cu \
    -Xlinker -lasan -Xlinker -lubsan ) > "$OUT/cpu_build_asan_lib.log" 2>&1 || { tail -5 "$OUT/cpu_build_asan_lib.log"; exit 1; }
set +x
PRE="$(gcc -print-file-name=libasan.so):$(gcc -print-file-name=libubsan.so)"
export ASAN_OPTIONS=detect_leaks=0:abort_on_error=0:halt_on_error=0:log_path=$OUT/cpu_asan_report
export UBSAN_OPTIONS=print_stacktrace=1:halt_on_error=0:log_path=$OUT/cpu_ubsan_report
rm -f "$OUT"/cpu_asan_report.* "$OUT"/cpu_ubsan_report.*
cd "$ROOT"
echo "== oracle + reference binary under ASan/UBSan (tests/test_oracle.py)"
LD_PRELOAD=$PRE TFPD_ORACLE_LIBDIR=$SAN python -m pytest tests/test_oracle.py -q -x -p no:cacheprovider 2>&1 | tail -3 | tee "$OUT/cpu_pytest_oracle.log"
echo "== product library host side under ASan/UBSan (tests/test_host.py: symbols, structs, argument checks)"
LD_PRELOAD=$PRE TFPOSE_DECODE_LIB=$ROOT/tools/ab/libtfpose_decode_asan.so python -m pytest tests/test_host.py -q -x -p no:cacheprovider 2>&1 | tail -3 | tee "$OUT/cpu_pytest_host.log"
echo "== sanitizer reports:"
ls "$OUT" | grep -E "cpu_(asan|ubsan)_report" || echo "none (no ASan / UBSan finding)"
