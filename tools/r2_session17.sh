#!/bin/bash
# 8-GPU e2e step with RLB_COMM_TRACE=1 (phase durations of scorer creation and of a round, rank 0)
N=${1:-8}; T=${2:-r02s}; O=gpurun_out; mkdir -p $O
RLB_COMM_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 \
  bench.py --gpus $N --steps 5 --warmup 3 --no-localization --no-extra > $O/${T}_trace_g${N}.json 2> $O/${T}_trace_g${N}.err; echo "rc=$?"
grep "rlb_comm trace" $O/${T}_trace_g${N}.err | tail -12
python -c "
import json;d=json.loads(open('$O/${T}_trace_g${N}.json').read().strip().splitlines()[-1]);print(d['value'],d['e2e']['value'],d['e2e']['breakdown_ms'])"
