#!/bin/bash
# multi-GPU session: the >= 2 GPU tests, then bench.py as the driver launches it (torchrun, N ranks)
N=${1:-8}; T=${2:-r02q}; O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8 > $O/${T}_g${N}_gpus.txt
timeout 300 python -m pytest tests/test_gpu_comm.py -x -q > $O/${T}_g${N}_pytest_comm.log 2>&1; echo "pytest comm rc=$?"; tail -2 $O/${T}_g${N}_pytest_comm.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --steps 10 --warmup 3 > $O/${T}_bench_c5_g${N}.json 2> $O/${T}_bench_c5_g${N}.err; echo "bench g$N rc=$?"
head -c 600 $O/${T}_bench_c5_g${N}.json; echo
