#!/bin/bash
# GPU-box call: the staged strip form of the bicubic resize: parity tests, then times against the other forms for
# strip heights 16 / 32 and a few CTA widths.  bash tools/gpu_round_check11.sh [prefix]
P=${1:-r02x}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_preproc.py -m gpu -x -q > gpurun_out/${P}_pytest_preproc.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/${P}_pytest_preproc.log
for ty in 32 16; do for th in 0 128 160 224 256; do
  echo "== TFPD_RESIZE_TY=$ty TFPD_RESIZE_THREADS=$th"
  TFPD_RESIZE_TY=$ty TFPD_RESIZE_THREADS=$th timeout 300 python tools/preproc_times.py 2>&1 | grep -E "staged|resize_cubic strip" | tee -a gpurun_out/${P}_preproc_times.log
done; done
timeout 300 python tools/preproc_times.py 64 720 1280 2>&1 | grep -E "resize_cubic" | tee -a gpurun_out/${P}_preproc_times.log
