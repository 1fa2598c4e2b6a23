#!/bin/bash
# the driver's flow on a 2-GPU box: reference arm alone and under torchrun, then N = 1, 2
O=gpurun_out; mkdir -p $O
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > $O/r29_ref_n2.json 2> $O/r29_ref_n2.err; echo "ref under torchrun rc=$?"; wc -l < $O/r29_ref_n2.json; head -c 250 $O/r29_ref_n2.json; echo
bash tools/r2_session8.sh r29
