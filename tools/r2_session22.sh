#!/bin/bash
# sorted two-speed tensor-core sweep (variant 12): parity (bounded), then bench --profile with and without it
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "sorted_two_speed" > gpurun_out/r22_pytest_tc2.log 2>&1; rc=$?; echo "pytest tc2 rc=$rc"; tail -30 gpurun_out/r22_pytest_tc2.log | cut -c1-300
for v in 1 0; do
  echo -n "[RLB_TC_SORTED=$v] "; RLB_TC_SORTED=$v timeout 200 python bench.py --profile --steps 5 --warmup 3 --no-localization 2>&1 | tail -1 | head -c 400; echo
done
echo -n "[no counts] "; timeout 200 python bench.py --profile --steps 5 --warmup 3 --no-localization --no-counts 2>&1 | tail -1 | head -c 400; echo
