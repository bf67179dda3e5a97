#!/bin/bash
# Race / memory checking of the single-GPU kernels.  One compute-sanitizer tool per
# gpurun call (B200_PROFILING.md): usage  scripts/sanitize.sh memcheck|racecheck|synccheck|initcheck
# Cross-GPU flag protocols are not visible to racecheck; they are covered by the stress loop in
# scripts/mgpu_check.py (--stress N: N back-to-back mixed collectives with randomised per-rank delays).
# NOTE: on the graft GPU pool compute-sanitizer is refused ("closed on this pool"); see DESIGN.md section 7 for
# what replaces it there.
set -euo pipefail
TOOL=${1:-memcheck}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "norm or rope or embed or sample or attn_decode" > gpurun_out/sanitize_plain.log 2>&1
compute-sanitizer --tool "$TOOL" --error-exitcode 1 --kernel-regex kns=norm_residual,kns=rope_kv,kns=embed_kernel,kns=sample_kernel \
  python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "norm or rope or embed or sample" > "gpurun_out/sanitize_$TOOL.log" 2>&1
tail -5 "gpurun_out/sanitize_$TOOL.log"
