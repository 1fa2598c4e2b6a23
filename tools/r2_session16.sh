#!/bin/bash
# ncu --set full of localize_cta_kernel over 592 and 2368 queries (one / four per resident CTA slot): DRAM traffic against
# the algorithmic input (queries x 2000 matches x 40 B), issue and pipe utilisation
O=gpurun_out; T=${1:-r02r}; mkdir -p $O
for Q in 592 2368; do
  timeout 100 python tools/time_localize.py $Q dev 2 2>&1 | tail -1
  timeout 900 ncu --set full --clock-control none -k regex:localize_cta_kernel -c 1 -f -o $O/${T}_localize_q$Q \
     python tools/time_localize.py $Q dev 1 > $O/${T}_ncu_localize_q$Q.log 2>&1; echo "ncu q=$Q rc=$?"
  python tools/ncu_summary.py $O/${T}_localize_q$Q.ncu-rep $O/${T}_ncu_localize_q$Q.csv
done
grep -h "dram__bytes\|gpu__time_duration\|issue_active\|fp64\|warps_active\|long_scoreboard\|wait_per" $O/${T}_ncu_localize_q*.csv
