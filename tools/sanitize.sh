#!/bin/bash
# compute-sanitizer over the hot path (run under gpurun): bash tools/sanitize.sh <tag>
# memcheck + racecheck + synccheck on smoke() (single-point kernels, 40 batched moves of 2 chains) and on the parity
# tests that reach the other kernel instantiations (swap, reservoir, tilted cells, k-sharded steps).
tag=${1:-rX}
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
SMOKE="python -c 'import __graft_entry__ as g; g.smoke()'"
TESTS="python -m pytest tests/test_gpu_parity.py -x -q -k 'swap_batched or reservoir or triclinic_strict or k_sharded_steps'"
for tool in memcheck racecheck synccheck; do
  for what in SMOKE TESTS; do
    log=gpurun_out/${tag}_sanitizer_${tool}_${what,,}.log
    eval timeout 360 $CS --tool $tool --print-limit 20 ${!what} > $log 2>&1
    echo "== $tool $what rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $log | tail -n 1) | $(grep -E 'passed|failed|smoke ok' $log | tail -n 1)"
  done
done | tee gpurun_out/${tag}_sanitizer_summary.txt
