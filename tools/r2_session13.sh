#!/bin/bash
# inlier pass A/B builds: EXTRA flags -> bench --profile line (inlier_ms = one K = 1 pass over N = 1M, L2 flushed)
for x in "$@"; do
  touch ransaclib_b200/csrc/score_kernels.cu
  make -s -C ransaclib_b200/csrc EXTRA="$x" 2>&1 | grep -A3 "inlier_range_kernelINS_5PoseF" | grep Used | sed 's/ptxas info    : //'
  echo -n "[$x] "; timeout 200 python bench.py --profile --steps 8 --warmup 3 --no-localization 2>/dev/null | head -c 300; echo
done
