#!/bin/bash
# timing probes of the tensor-core sweep (results are wrong by construction): which resource bounds it?
for x in "-DRLB_TC_PROBE_NOMUFU" "-DRLB_TC_PROBE_NOLDTM" "-DRLB_TC_PROBE_NOMUFU -DRLB_TC_PROBE_NOLDTM" "-DRLB_TC_PIPE=0 -DRLB_TC_PROBE_NOMUFU" "-DRLB_TC_PIPE=0"; do
  touch ransaclib_b200/csrc/sweep_tc_kernel.cu
  make -s -C ransaclib_b200/csrc EXTRA="$x" > /dev/null 2>&1
  echo -n "[$x] "; timeout 200 python bench.py --profile --steps 5 --warmup 3 --no-localization 2>/dev/null | head -c 250; echo
done
