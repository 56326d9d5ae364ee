#!/bin/bash
# compute-sanitizer over the unit-size GPU tests (memcheck and racecheck), run on the GPU box:
#   bash profiles/run_sanitizer.sh            -> gpurun_out/r02_sanitizer_{memcheck,racecheck}.txt
# The selection keeps every kernel family (layout, GP, the three solve kernels, advance, ingest, output) at
# sizes the tools finish in minutes; pytest runs inside the tool, so every kernel the tests launch is checked.
OUT=${1:-gpurun_out}
SEL='tests/test_gpu_layout.py tests/test_gpu_advance.py tests/test_gpu_read_small.py tests/test_gpu_assimilate_iter.py tests/test_zz_no_alloc.py'
SOLVE='tests/test_gpu_solve.py -k "(all_state_sizes and (3- or 10- or 13- or 16-)) or pair_kernel or mapper_linear or non_positive"'
GP='tests/test_gpu_gp.py -k "predict_mean_jacobian and (4-120 or 10-250) or underflow or linearize_bands_mask and (127 or 4095)"'
for tool in memcheck racecheck; do
  log=$OUT/r02_sanitizer_$tool.txt
  : > $log
  for part in "$SEL" "$SOLVE" "$GP"; do
    echo "### compute-sanitizer --tool $tool python -m pytest $part" >> $log
    eval timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 5 \
      python -m pytest $part -m gpu -x -q -p no:cacheprovider 2>&1 | grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|error|Error|hazard" | tail -12 >> $log
    echo "exit status: ${PIPESTATUS[0]}" >> $log
  done
done
for tool in memcheck racecheck; do echo "== $tool"; tail -40 $OUT/r02_sanitizer_$tool.txt; done
