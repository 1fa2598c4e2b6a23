#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r26_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r26_pytest.log
timeout 200 python bench.py --profile --steps 5 --warmup 3 --no-localization 2>/dev/null | head -c 300; echo
