#!/bin/bash
# localize_cta_kernel A/B: shared-memory staging of the f32 copy, non-allocating point loads, CTAs per SM
O=gpurun_out; T=r02k; mkdir -p $O
ab() {
  touch ransaclib_b200/csrc/localize_kernels.cu
  make -s -C ransaclib_b200/csrc EXTRA="$1" 2>&1 | grep -A2 "localize_cta_kernelENS" | grep Used | sed 's/ptxas info    : //'
  for Q in 512 4096; do echo -n "[$1] "; timeout 100 python tools/time_localize.py $Q dev 3 2>&1 | tail -1; done
}
( ab ""
  ab "-DRLB_LOC_CTA_SMEM=0"
  ab "-DRLB_LOC_STREAM=1"
  ab "-DRLB_LOC_CTA_SMEM=0 -DRLB_LOC_STREAM=1"
  ab "-DRLB_LOC_CTA_SMEM=0 -DRLB_LOC_STREAM=1 -DRLB_LOC_CTA_MINB=3"
  ab "-DRLB_LOC_CTA_SMEM=0 -DRLB_LOC_STREAM=1 -DRLB_LOC_CTA_MINB=5"
  ab "-DRLB_LOC_CTA_SMEM=0 -DRLB_LOC_STREAM=1 -DRLB_LOC_CTA_MINB=6"
) 2>&1 | tee $O/${T}_ab.txt
touch ransaclib_b200/csrc/localize_kernels.cu; make -s -C ransaclib_b200/csrc EXTRA="-DRLB_LOC_CTA_SMEM=0 -DRLB_LOC_STREAM=1" > /dev/null 2>&1
(timeout 100 python tools/prof_localize.py 1; timeout 100 python tools/prof_localize.py 592) 2>&1 | tee $O/${T}_phases.txt
timeout 200 python tools/timeline_localize.py 4096 1 2>&1 | tee $O/${T}_timeline.txt
