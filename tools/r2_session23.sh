#!/bin/bash
# ncu --set full of the sorted two-speed sweep's main kernel (K = 16384 to keep the replays short)
O=gpurun_out; mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweep_pose_tc2_kernel -c 2 -f -o $O/r23_tc2 \
  python bench.py --profile --steps 1 --warmup 1 --k-override 16384 --no-localization > $O/r23_ncu_tc2.log 2>&1; echo "ncu tc2 rc=$?"
python tools/ncu_summary.py $O/r23_tc2.ncu-rep $O/r23_ncu_tc2.csv
grep -i "dram__bytes_read.sum,\|dram__bytes_write.sum,\|smsp__inst_executed.sum,\|issue_active.avg.pct\|gpu__time_duration.sum\|pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed\|inst_executed_pipe_xu.avg.pct\|pipe_alu_cycles\|pipe_tensor\|stalled.*per_issue_active\|lts__t_sector_hit_rate\|branch_targets_threads_divergent\|sm__warps_active.avg.pct" $O/r23_ncu_tc2.csv | head -60
