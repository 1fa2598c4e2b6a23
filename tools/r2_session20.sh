#!/bin/bash
# decoupled tensor-core sweep (sweep_pose_tcp_kernel): parity first (bounded), then the bench's profile line
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tensor_core or tc" > gpurun_out/r20_pytest_tc.log 2>&1; rc=$?; echo "pytest tc rc=$rc"; tail -5 gpurun_out/r20_pytest_tc.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_end_to_end.py tests/test_gpu_comm.py -x -q -m gpu > gpurun_out/r20_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r20_pytest.log
for v in 1 2; do
  timeout 200 python bench.py --profile --steps 8 --warmup 3 --no-localization 2>/dev/null | head -c 400; echo
done
timeout 200 python bench.py --profile --steps 8 --warmup 3 --no-localization --no-counts 2>/dev/null | head -c 400; echo
