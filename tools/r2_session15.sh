#!/bin/bash
# f32 tensor-core sweep A/B builds: EXTRA flags -> bench --profile line (C5, sweep_ms)
for x in "$@"; do
  touch ransaclib_b200/csrc/sweep_tc_kernel.cu
  make -s -C ransaclib_b200/csrc EXTRA="$x" 2>&1 | grep -A2 "sweep_pose_tc_kernelILb1" | grep Used | sed 's/ptxas info    : //'
  echo -n "[$x] "; timeout 200 python bench.py --profile --steps 5 --warmup 3 --no-localization 2>/dev/null | head -c 300; echo
  echo -n "[$x --no-counts] "; timeout 200 python bench.py --profile --steps 5 --warmup 3 --no-localization --no-counts 2>/dev/null | head -c 300; echo
done
