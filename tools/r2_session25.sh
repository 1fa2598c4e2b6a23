#!/bin/bash
# per-phase clock64 stamps of one CTA of the sorted two-speed sweep (hypothesis block 300 = an unflagged one)
for x in "-DRLB_TC2_STAMPS" "-DRLB_TC2_STAMPS -DRLB_TC2_PROBE_NOEPI"; do
  touch ransaclib_b200/csrc/sweep_tc_kernel.cu
  make -s -C ransaclib_b200/csrc EXTRA="$x" > /dev/null 2>&1
  echo "[$x]"; timeout 200 python bench.py --profile --steps 1 --warmup 1 --no-localization 2>&1 | grep -E "^producer|^group" | tail -3
done
