#!/bin/bash
# full GPU test suite, smoke, bench N=1 (default line), then N=2 under torchrun (comm path)
O=gpurun_out; T=${1:-r02e}; mkdir -p $O
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv > $O/${T}_gpu.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${T}_pytest.log
timeout 200 python __graft_entry__.py smoke > $O/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/${T}_smoke.log
timeout 600 python bench.py > $O/${T}_bench_n1.json 2> $O/${T}_bench_n1.err; echo "bench n1 rc=$?"; tail -3 $O/${T}_bench_n1.err; head -c 1500 $O/${T}_bench_n1.json; echo
NG=$(nvidia-smi -L | wc -l)
if [ "$NG" -ge 2 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > $O/${T}_bench_n2.json 2> $O/${T}_bench_n2.err; echo "bench n2 rc=$?"; tail -5 $O/${T}_bench_n2.err; head -c 1500 $O/${T}_bench_n2.json; echo
fi
