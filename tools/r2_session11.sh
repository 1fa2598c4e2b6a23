#!/bin/bash
# ncu --set full of the filtered f64 sweep (sweep_pose_tcx_kernel), K = 16384 of BASELINE configs[4]
O=gpurun_out; T=${1:-r02n}; mkdir -p $O
timeout 300 python bench.py --precision f64 --profile --steps 1 --warmup 1 --k-override 16384 > $O/${T}_tcx_plain.json 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweep_pose_tcx -c 1 -f -o $O/${T}_tcx \
   python bench.py --precision f64 --profile --steps 1 --warmup 1 --k-override 16384 > $O/${T}_ncu_tcx.log 2>&1; echo "ncu tcx rc=$?"
python tools/ncu_summary.py $O/${T}_tcx.ncu-rep $O/${T}_ncu_tcx.csv; cat $O/${T}_ncu_tcx.csv | head -50
