#!/bin/bash
# Runs ON THE GPU BOX (gpurun): compute-sanitizer over the tiny-shape GPU tests (memcheck / racecheck targets).
# The sanitizer slows kernels 10-100x, so only small fixtures are selected; each tool has its own time limit.
# Summaries -> gpurun_out/sanitize/{memcheck,racecheck}.log (copied to profiles/r2_sanitizer.txt).
set -u
O=gpurun_out/sanitize; mkdir -p $O
SEL='cfg1 or edge_cases or remove_mean or sampler_step or diffuse_as_batch or fc_edges or cutoff or graph_index or noalloc or selftest or radam'
FILES="tests/test_gpu_forward.py tests/test_gpu_backward.py tests/test_gpu_tc.py tests/test_gpu_optim_ckpt.py tests/test_zz_no_alloc.py"
for TOOL in memcheck racecheck; do
  timeout 420 compute-sanitizer --tool $TOOL --error-exitcode 9 --launch-timeout 120 --print-limit 20 \
      python -m pytest $FILES -m gpu -q -x -k "$SEL" -p no:cacheprovider --timeout 400 > $O/$TOOL.log 2>&1
  echo "$TOOL rc=$?" | tee -a $O/summary.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Invalid|hazard" $O/$TOOL.log | tail -8 | tee -a $O/summary.txt
done
