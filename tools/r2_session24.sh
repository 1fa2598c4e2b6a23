#!/bin/bash
# timing probes of the sorted two-speed sweep (results wrong by construction)
for x in "-DRLB_TC2_PROBE_PZ" "-DRLB_TC2_PROBE_NOSLOW" "-DRLB_TC2_PROBE_NOEPI" ""; do
  touch ransaclib_b200/csrc/sweep_tc_kernel.cu
  make -s -C ransaclib_b200/csrc EXTRA="$x" > /dev/null 2>&1
  echo -n "[$x] "; timeout 200 python bench.py --profile --steps 4 --warmup 2 --no-localization 2>/dev/null | head -c 250; echo
done
