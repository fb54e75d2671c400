#!/bin/bash
# Memory-safety runs WITHOUT compute-sanitizer (closed on this GPU pool: profiles/r02_sanitizers/README.md):
# the GPU test-suite against the CHECKED build of the library (make -C tf-pose-estimation_b200/csrc checked):
#   - TFPD_CHK assertions on the data-dependent indices (K1 slot pokes and peak coordinates, K2 gather coordinates and
#     table indices), guard bands between the shared-memory sub-arrays of K1 and K3 verified when a role finishes,
#   - TFPD_CHECKED=1: a 256-byte guard band behind every workspace array in global memory, verified by every fetch,
# plus two NEGATIVE CONTROLS showing that the mechanisms fire (a poked workspace guard fails the fetch; a deliberate
# one-past-the-end write in the assembly kernel trips the shared-memory guard and traps).
# Output: gpurun_out/r02_sanitizers/checked_*.log
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/gpurun_out/r02_sanitizers
LIB=$ROOT/tools/ab/libtfpose_decode_checked.so
mkdir -p "$OUT"; cd "$ROOT"
[ -f "$LIB" ] || { echo "checked library missing: make -C tf-pose-estimation_b200/csrc checked"; exit 1; }
TFPD_CHECKED=1 TFPOSE_DECODE_LIB=$LIB python -m pytest tests -m gpu -q -p no:cacheprovider > "$OUT/checked_pytest.log" 2>&1
echo "checked build, GPU suite: exit $? | $(tail -1 "$OUT/checked_pytest.log") | TFPD_CHK messages: $(grep -c 'TFPD_CHK failed' "$OUT/checked_pytest.log")"
TFPD_CHECKED=1 TFPOSE_DECODE_LIB=$LIB python tools/sanitize_target.py > "$OUT/checked_target.log" 2>&1
echo "checked build, sanitize_target: exit $? | $(tail -1 "$OUT/checked_target.log")"
# negative control 1: a poked workspace guard band must fail the next fetch
TFPD_CHECKED=1 TFPOSE_DECODE_LIB=$LIB python - > "$OUT/checked_negative_workspace.log" 2>&1 <<'PY'
import ctypes as C, sys, torch
sys.path.insert(0, ".")
import tfpose_b200 as tp
from tfpose_b200 import synth, _lib
dec = tp.PoseDecoder(max_batch=2, max_upsample=(184, 216))
x = torch.from_numpy(synth.synth_batch(2, 46, 54, config_id=2)).cuda()
dec.decode(x, (184, 216))
print("clean decode: ok")
poke = _lib.lib().tfpd_debug_poke_guard
poke.argtypes, poke.restype = [C.c_void_p, C.c_int], C.c_int
assert poke(dec._ctx, 3) == 0
try:
    dec.decode(x, (184, 216))
    print("NEGATIVE CONTROL FAILED: poked guard not noticed")
    sys.exit(1)
except tp.DecodeError as e:
    print("poked guard noticed:", e)
PY
echo "negative control (workspace guard): exit $? | $(tail -1 "$OUT/checked_negative_workspace.log" | cut -c1-160)"
# negative control 2: a one-past-the-end write in the assembly kernel must trip the shared-memory guard (the kernel traps)
TFPD_CHECKED=1 TFPD_CHECKED_TRIP=1 TFPOSE_DECODE_LIB=$LIB timeout 120 python tools/sanitize_target.py > "$OUT/checked_negative_smem.log" 2>&1
rc=$?
echo "negative control (shared-memory guard, expected to fail): exit $rc | $(grep -m1 'TFPD_CHK failed' "$OUT/checked_negative_smem.log" | cut -c1-160)"
[ $rc -ne 0 ] && grep -q "TFPD_CHK failed" "$OUT/checked_negative_smem.log" && echo "negative controls OK"
