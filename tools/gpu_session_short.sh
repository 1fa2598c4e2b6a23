#!/bin/bash
# Short gpurun session (no ncu --set full captures): gpu tests, smoke, the bench lines, launch list, line LO-MSAC
# timing and the configs[0..2] end-to-end table. Usage: gpurun --timeout 700 -- 'bash tools/gpu_session_short.sh [tag]'
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${TAG}_gpu.txt 2>&1
nproc >> $O/${TAG}_gpu.txt
timeout 700 python -m pytest tests -m gpu -x -q --durations=15 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
tail -3 $O/${TAG}_pytest.log
timeout 120 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/${TAG}_smoke.log
timeout 200 python tests/e2e_configs.py > $O/${TAG}_e2e_configs.json 2> $O/${TAG}_e2e_configs.err; echo "e2e configs rc=$?"
timeout 100 python tools/time_line_lomsac.py 592 10000 0.5 > $O/${TAG}_line_lomsac.json 2> $O/${TAG}_line_lomsac.err; echo "line lomsac rc=$?"
timeout 300 python bench.py > $O/${TAG}_bench_c5.json 2> $O/${TAG}_bench_c5.err; echo "bench c5 rc=$?"
timeout 300 python bench.py --workload c4 --steps 5 --warmup 2 > $O/${TAG}_bench_c4.json 2> $O/${TAG}_bench_c4.err; echo "bench c4 rc=$?"
timeout 200 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_c5_ref.json 2>&1; echo "bench ref rc=$?"
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --profile --steps 2 --warmup 1 > $O/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
head -c 500 $O/${TAG}_bench_c5.json; echo
head -c 500 $O/${TAG}_bench_c4.json; echo
