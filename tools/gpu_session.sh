#!/bin/bash
# One gpurun session: gpu tests, bench lines, ncu launch list + one full capture of the sweep.
# Usage: gpurun --timeout 1500 -- 'bash tools/gpu_session.sh [tag]'
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${TAG}_gpu.txt 2>&1
nproc >> $O/${TAG}_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
tail -3 $O/${TAG}_pytest.log
timeout 120 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/${TAG}_smoke.log
timeout 400 python bench.py > $O/${TAG}_bench_c5.json 2> $O/${TAG}_bench_c5.err; echo "bench c5 rc=$?"
timeout 200 python bench.py --no-localization --no-counts > $O/${TAG}_bench_c5_nocounts.json 2> $O/${TAG}_bench_c5_nocounts.err; echo "bench c5 no-counts rc=$?"
RLB_SWEEP_VARIANT=3 timeout 200 python bench.py --no-localization > $O/${TAG}_bench_c5_ffma2.json 2> $O/${TAG}_bench_c5_ffma2.err; echo "bench c5 ffma2 rc=$?"
timeout 300 python bench.py --precision f64 --steps 3 > $O/${TAG}_bench_c5_f64.json 2> $O/${TAG}_bench_c5_f64.err; echo "bench c5 f64 rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_c5_ref.json 2>&1; echo "bench ref rc=$?"
timeout 400 python bench.py --workload c4 --steps 5 --warmup 2 > $O/${TAG}_bench_c4.json 2> $O/${TAG}_bench_c4.err; echo "bench c4 rc=$?"
timeout 300 python bench.py --workload c4 --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_c4_ref.json 2>&1; echo "bench c4 ref rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --profile --steps 2 --warmup 1 > $O/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweep -c 2 -f -o $O/${TAG}_sweep \
  python bench.py --profile --steps 1 --warmup 1 > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:inlier -c 2 -f -o $O/${TAG}_inlier \
  python bench.py --profile --steps 1 --warmup 1 > $O/${TAG}_ncu_inl.log 2>&1; echo "ncu inlier rc=$?"
# device-pointer entry point only: under ncu the pipelined host API would wait for copies that never start
timeout 400 ncu --set full --clock-control none --import-source on -k regex:localize -c 1 -f -o $O/${TAG}_localize \
  python tools/time_localize.py 2368 dev 1 > $O/${TAG}_ncu_localize.log 2>&1; echo "ncu localize rc=$?"
# device-resident line LO-MSAC: latency / throughput / phase counters, the configs[0..2] end-to-end table, ncu of the
# first launch (one problem = one CTA: the latency view)
timeout 200 python tools/time_line_lomsac.py 592 10000 0.5 > $O/${TAG}_line_lomsac.json 2> $O/${TAG}_line_lomsac.err; echo "line lomsac rc=$?"
timeout 400 python tests/e2e_configs.py > $O/${TAG}_e2e_configs.json 2> $O/${TAG}_e2e_configs.err; echo "e2e configs rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:line_lomsac -c 1 -f -o $O/${TAG}_line_lomsac \
  python tools/time_line_lomsac.py 8 10000 0.5 > $O/${TAG}_ncu_line_lomsac.log 2>&1; echo "ncu line lomsac rc=$?"
head -c 600 $O/${TAG}_bench_c5.json; echo
head -c 600 $O/${TAG}_bench_c4.json; echo
