#!/bin/bash
O=gpurun_out; T=r02d; mkdir -p $O
(timeout 200 python tools/timeline_localize.py 4096 1; timeout 200 python tools/timeline_localize.py 512 1; timeout 200 python tools/timeline_localize.py 4096 0) 2>&1 | tee $O/${T}_timeline.txt
