#!/bin/bash
# A/B of sweep_tc_kernel build switches on the GPU box: EXTRA flags per variant (results of
# experiment builds may be wrong on purpose; the last build is left in place, so rebuild afterwards).
export RLB_SWEEP_VARIANT=9
for v in "$@"; do
  touch ransaclib_b200/csrc/sweep_tc_kernel.cu
  make -s -C ransaclib_b200/csrc EXTRA="$v" 2>&1 | grep -A2 "sweep_pose_tc_kernel" | grep Used
  echo "variant: $v"
  timeout 60 python bench.py --steps 5 --warmup 3 --no-localization 2>&1 | tail -1 | cut -c1-140
done
