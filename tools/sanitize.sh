#!/bin/bash
# compute-sanitizer runs over the small cases of every hand-written kernel.  One tool per gpurun call
# (B200_PROFILING.md): tools/sanitize.sh memcheck | racecheck | synccheck
# NOTE (rounds 1 and 2): on the shared B200 pool `compute-sanitizer` answers "closed on this pool and stays closed"
# (exit 86; profiles/sanitizer_r2.md has the message and the substitutes: bounded waits, tail shapes against fp32
# oracles, bitwise repeatability, tools/comm_stress.py).
set -u
cd "$(dirname "$0")/.."
TOOL=${1:-memcheck}
for cmd in "tools/gemm_check.py nt_small" "tools/gemm_check.py nt_tail" "tools/gemm_check.py tn_small" "tools/gemm_check.py swiglu_small" \
           "tools/gemm_check.py rope_small" "tools/attn_check.py s256_hd64" "tools/attn_check.py s384_hd128_mha"; do
  echo "== $TOOL: $cmd"
  timeout 300 compute-sanitizer --tool $TOOL --print-limit 5 python $cmd 2>&1 | grep -E "ERROR SUMMARY|Invalid|Race|hazard|\"ok\"" | cut -c1-200 | head -8
done
echo "== $TOOL: memory-bound kernels (pytest subset)"
timeout 600 compute-sanitizer --tool $TOOL --print-limit 5 python -m pytest tests/test_native_ops.py -q -m gpu -k "rmsnorm or rope_inplace or swiglu or embedding or grad_norm" 2>&1 | grep -E "ERROR SUMMARY|Invalid|passed|failed" | cut -c1-200 | head -8
