#!/bin/bash
O=gpurun_out; T=${1:-r02c}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_localize_batch.py tests/test_gpu_comm.py -x -q > $O/${T}_pytest_loc.log 2>&1; echo "pytest loc+comm rc=$?"; tail -15 $O/${T}_pytest_loc.log
for mode in 1 0; do
  for Q in 512 4096; do
    echo -n "mode $mode dev: "; RLB_LOC_MODE=$mode timeout 120 python tools/time_localize.py $Q dev 3 2>&1 | tail -1
  done
done 2>&1 | tee $O/${T}_time_localize.txt
echo -n "mode 1 host 4096: "; timeout 120 python tools/time_localize.py 4096 host 3 2>&1 | tail -1 | tee -a $O/${T}_time_localize.txt
(timeout 100 python tools/prof_localize.py 1; timeout 100 python tools/prof_localize.py 592) 2>&1 | tee $O/${T}_phases.txt
