#!/bin/bash
# Round-2 evidence session on one B200: gpu tests, smoke, the bench lines, launch list, ncu --set full summaries of
# the kernels that changed this round, configs[0..2] end-to-end table. Usage: gpurun --timeout 1500 -- 'bash tools/r2_session_final.sh r02z'
T=${1:-r02z}; O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${T}_gpu.txt 2>&1; nproc >> $O/${T}_gpu.txt
timeout 700 python -m pytest tests -m gpu -x -q --durations=10 > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${T}_pytest.log; tail -3 $O/${T}_pytest.log
timeout 120 python __graft_entry__.py smoke > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/${T}_smoke.log
timeout 400 python bench.py > $O/${T}_bench_c5.json 2> $O/${T}_bench_c5.err; echo "bench c5 rc=$?"
timeout 300 python bench.py --workload c4 --steps 5 --warmup 3 > $O/${T}_bench_c4.json 2> $O/${T}_bench_c4.err; echo "bench c4 rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_c5_ref.json 2>&1; echo "bench ref rc=$?"
timeout 200 python tests/e2e_configs.py > $O/${T}_e2e_configs.json 2> $O/${T}_e2e_configs.err; echo "e2e configs rc=$?"
timeout 100 python tools/time_line_lomsac.py 592 10000 0.5 > $O/${T}_line_lomsac.json 2> $O/${T}_line_lomsac.err; echo "line lomsac rc=$?"
# launch list of one default step (f32) and one f64 step
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches.csv \
  python bench.py --profile --steps 2 --warmup 1 --no-localization > $O/${T}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${T}_launches_f64.csv \
  python bench.py --precision f64 --profile --steps 2 --warmup 1 --no-localization > $O/${T}_ncu_launches_f64.log 2>&1; echo "ncu launches f64 rc=$?"
# ncu --set full: the filtered f64 sweep's sparse kernel, the dense-list f64 sweep, the inlier pass (K = 16384 to keep the replays short)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweep_pose_tcx_kernel -c 2 -f -o $O/${T}_tcx \
  python bench.py --precision f64 --profile --steps 1 --warmup 1 --k-override 16384 --no-localization > $O/${T}_ncu_tcx.log 2>&1; echo "ncu tcx rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:inlier_range_kernel -c 1 -f -o $O/${T}_inlier \
  python bench.py --profile --steps 1 --warmup 1 --k-override 16384 --no-localization > $O/${T}_ncu_inlier.log 2>&1; echo "ncu inlier rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweep_pose_tc_kernel -c 1 -f -o $O/${T}_tc \
  python bench.py --profile --steps 1 --warmup 1 --k-override 16384 --no-localization > $O/${T}_ncu_tc.log 2>&1; echo "ncu tc rc=$?"
python tools/ncu_summary.py $O/${T}_tc.ncu-rep $O/${T}_ncu_sweep_tc.csv
python tools/ncu_summary.py $O/${T}_tcx.ncu-rep $O/${T}_ncu_tcx.csv
python tools/ncu_summary.py $O/${T}_inlier.ncu-rep $O/${T}_ncu_inlier_range.csv
head -c 400 $O/${T}_bench_c5.json; echo
