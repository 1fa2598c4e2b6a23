#!/bin/bash
# the no-alloc test alone (fresh process), the whole GPU suite (it runs last there), then the scaling legs the box has
O=gpurun_out; T=${1:-r02j}; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > $O/${T}_pytest_noalloc_alone.log 2>&1; echo "no-alloc alone rc=$?"; tail -4 $O/${T}_pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest all rc=$?"; tail -4 $O/${T}_pytest.log
bash tools/r2_session8.sh $T
