#!/bin/bash
# A/B of localize_kernel build switches on the GPU box: EXTRA flags per variant.
for v in "$@"; do
  touch ransaclib_b200/csrc/localize_kernels.cu
  make -s -C ransaclib_b200/csrc EXTRA="$v" 2>&1 | grep -A2 "localize_cta_kernelENS" | tail -1
  echo "variant: $v"
  timeout 100 python tools/time_localize.py 4096 dev 3 2>&1 | tail -1
done
