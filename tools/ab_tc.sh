#!/bin/bash
# A/B of sweep_tc_kernel build switches on the GPU box: EXTRA flags per variant (results of
# experiment builds may be wrong on purpose; the last build is left in place, so rebuild afterwards).
# BENCH_FLAGS: extra bench.py flags (e.g. --counts).
export RLB_SWEEP_VARIANT=9
for v in "$@"; do
  touch ransaclib_b200/csrc/sweep_tc_kernel.cu
  make -s -C ransaclib_b200/csrc EXTRA="$v" 2>&1 | grep -A2 "sweep_pose_tc_kernel" | grep Used | tr '\n' ' '
  echo; echo "variant: $v"
  timeout 60 python bench.py --steps 5 --warmup 3 --no-localization $BENCH_FLAGS 2>&1 | tail -1 | cut -c1-110
done
