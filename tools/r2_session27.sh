#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r27_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r27_pytest.log
python tools/lat_probe.py 2>&1 | tail -3
