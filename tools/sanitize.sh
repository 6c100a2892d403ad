#!/usr/bin/env bash
# compute-sanitizer over the single-kernel drivers of tools/ncu_target.py (one GPU).  SURVEY.md §5.2: memcheck for every
# kernel; racecheck / synccheck for the ones that use shared memory + mbarriers.  Slow (10-100x): run a few targets at
# a time.  NOTE: on the round-1 GPU pool compute-sanitizer is administratively closed (every run returns rc 86 with a notice), so
# no sanitizer logs are committed; numerics tests against the fp32 oracle and the FDT_DEBUG_POISON mode stand in.
# usage: tools/sanitize.sh [tool] [target ...]     tool = memcheck | racecheck | synccheck | initcheck
set -u
tool=${1:-memcheck}; shift || true
targets=${*:-"chan_affine chan_reduce lin_prep ngd_axis ngd_eig wprep conv1x1 conv3x3 wgrad ngd_tc attn"}
mkdir -p gpurun_out
for t in $targets; do
  timeout 300 compute-sanitizer --tool "$tool" --error-exitcode 1 python tools/ncu_target.py "$t" > gpurun_out/san_${tool}_$t.log 2>&1
  rc=$?
  echo "$tool $t rc=$rc $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/san_${tool}_$t.log | tail -1)"
done
