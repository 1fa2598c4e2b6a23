#!/bin/bash
# Round 2, session 1: the reworked batched localization (stream-ordered chunk launches, one query per CTA).
O=gpurun_out
T=r02a
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${T}_gpu.txt 2>&1; nproc >> $O/${T}_gpu.txt
timeout 600 python -m pytest tests/test_gpu_localize_batch.py -x -q > $O/${T}_pytest_loc.log 2>&1; echo "pytest loc rc=$?"; tail -5 $O/${T}_pytest_loc.log
for mode in 1 0; do
  for Q in 512 4096; do
    echo "== mode $mode Q $Q dev"; RLB_LOC_MODE=$mode timeout 120 python tools/time_localize.py $Q dev 3 2>&1 | tail -2
    echo "== mode $mode Q $Q host"; RLB_LOC_MODE=$mode timeout 120 python tools/time_localize.py $Q host 3 2>&1 | tail -2
  done
done 2>&1 | tee $O/${T}_time_localize.txt
echo "== phases, 1 query alone / 592 queries (mode 1)"; 
(timeout 100 python tools/prof_localize.py 1; timeout 100 python tools/prof_localize.py 592; RLB_LOC_MODE=0 timeout 100 python tools/prof_localize.py 1) 2>&1 | tee $O/${T}_phases.txt
echo "== A/B LM state in shared memory (mode 1)"
bash tools/ab_localize.sh "-DRLB_LM_SMEM=1" "-DRLB_LOC_CTA_MINB=3" "" 2>&1 | tee $O/${T}_ab.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest all rc=$?"; tail -5 $O/${T}_pytest.log
