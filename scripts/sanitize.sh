#!/bin/bash
# compute-sanitizer over the single-GPU kernel tests.  One tool per gpurun call:
#   gpurun --timeout 900 -- 'bash scripts/sanitize.sh memcheck'      (or racecheck / synccheck / initcheck)
# Cross-rank flag protocols are not visible to the sanitizer: they are covered by
# tests/test_collectives_gpu.py::test_emulated_collectives_under_random_rank_skew (ctx.skew_ns).
TOOL=${1:-memcheck}
mkdir -p gpurun_out
# small shapes only: the sanitizer slows kernels 10-100x
timeout 800 compute-sanitizer --tool "$TOOL" --error-exitcode 3 --log-file gpurun_out/sanitizer_${TOOL}.log \
    python -m pytest tests/test_kernels_gpu.py tests/test_attention_gpu.py tests/test_conv_gpu.py -m gpu -q -x \
    -k "adamw or sgd or cross_entropy or mse or layer_norm or unscale or sumsq or 5-2-17-32 or 16-6-64-10 or 16-16-1-8 or 19-23-4-24" \
    > gpurun_out/sanitizer_${TOOL}_pytest.log 2>&1
rc=$?
tail -3 gpurun_out/sanitizer_${TOOL}_pytest.log
grep -E "ERROR SUMMARY|Invalid|hazard" gpurun_out/sanitizer_${TOOL}.log | tail -5
exit $rc
