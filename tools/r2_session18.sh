#!/bin/bash
# inlier pass: one-CTA-per-SM bulk-TMA ring (default) vs the register-prefetch kernel (RLB_INLIER_TMA=0)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py tests/test_custom_functor.py tests/test_gpu_end_to_end.py -x -q -m gpu > gpurun_out/r18_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r18_pytest.log
for v in 1 0 1 0; do
  echo -n "[RLB_INLIER_TMA=$v] "; RLB_INLIER_TMA=$v timeout 200 python bench.py --profile --steps 8 --warmup 3 --no-localization 2>/dev/null | head -c 400; echo
done
