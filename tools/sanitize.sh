#!/bin/bash
# compute-sanitizer over a small-shape subset of the GPU parity tests: every product kernel family is launched at
# least once (group quant incl. v2 / masked layouts, per-token quant, every product GEMM variant, forced stream-K,
# swap-AB decode kernel in all three schedules, fused-quant epilogue, grouped GEMM, ragged K/N), at sizes the
# sanitizer finishes in minutes.  ONE tool per invocation (one per gpurun call):
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck
# Writes gpurun_out/sanitizer_<tool>.log (full) and prints the summary lines; tools/sanitize_summary.py folds the
# logs into profiles/sanitizer_r02.txt.
TOOL=${1:-memcheck}
mkdir -p gpurun_out
SUBSET=(
  "tests/test_quant_gpu.py::test_golden_fp8"
  "tests/test_quant_gpu.py::test_golden_int8"
  "tests/test_quant_gpu.py::test_golden_ue8m0_row_major_bytes"
  "tests/test_quant_gpu.py::test_ue8m0_packed_column_major"
  "tests/test_quant_gpu.py::test_zero_rows_and_empty"
  "tests/test_quant_gpu.py::test_misaligned_views_use_generic_kernel"
  "tests/test_quant_gpu.py::test_fused_silu_mul_v2"
  "tests/test_quant_gpu.py::test_v2_masked_moe_layout"
  "tests/test_quant_gpu.py::test_v2_column_major_packed_ue8m0"
  "tests/test_token_quant_gpu.py::test_golden"
  "tests/test_token_quant_gpu.py::test_shapes_signed"
  "tests/test_gemm_gpu.py::test_golden"
  "tests/test_gemm_gpu.py::test_every_variant"
  "tests/test_gemm_gpu.py::test_stream_k_matches_whole_tiles_and_oracle"
  "tests/test_gemm_gpu.py::test_decode_kernel"
  "tests/test_gemm_gpu.py::test_ragged_k_and_n"
  "tests/test_gemm_gpu.py::test_multi_destination_epilogue"
  "tests/test_gemm_quant_gpu.py::test_fused_equals_separate_ops"
  "tests/test_gemm_quant_gpu.py::test_fused_silu_mul_equals_separate_ops"
  "tests/test_grouped_gpu.py::test_golden"
  "tests/test_grouped_gpu.py::test_ragged_and_empty_experts"
)
# small parametrisations only (ids are "M-N-K[-...]"): anything with a dimension >= 4096 or M >= 512 is left to the
# plain GPU suite
KEEP='not 7168 and not 24576 and not 4096 and not 8192 and not 16384 and not 16392 and not 40960 and not 2000 and not 1024-1024 and not 512-2112 and not 1000- and not 513- and not 300-640 and not 300-1000'
OPTS="--tool $TOOL --error-exitcode 9 --print-limit 40 --launch-timeout 0"
if [ "$TOOL" = "memcheck" ]; then OPTS="$OPTS --leak-check no"; fi
SGLK_SANITIZER=1 compute-sanitizer $OPTS python -m pytest -q -x -p no:cacheprovider -m gpu "${SUBSET[@]}" -k "$KEEP" \
  > gpurun_out/sanitizer_$TOOL.log 2>&1
RC=$?
echo "exit code $RC" >> gpurun_out/sanitizer_$TOOL.log
grep -E "passed|failed|ERROR SUMMARY|RACECHECK SUMMARY|exit code|Error:|hazard" gpurun_out/sanitizer_$TOOL.log | sort | uniq -c | sort -rn | head -30
