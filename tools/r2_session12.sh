#!/bin/bash
# tcx A/B builds: EXTRA flags -> f64 bench line
for x in "$@"; do
  touch ransaclib_b200/csrc/sweep_tc_kernel.cu
  make -s -C ransaclib_b200/csrc EXTRA="$x" 2>&1 | grep -A2 "tcx" | grep Used | sed 's/ptxas info    : //'
  echo -n "[$x] "; timeout 200 python bench.py --precision f64 --profile --steps 3 --warmup 2 2>/dev/null | head -c 300; echo
done
