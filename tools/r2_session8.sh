#!/bin/bash
# scaling run as the driver does it: reference arm, then N = 1, 2, 4, 8 (as many as the box has)
O=gpurun_out; T=${1:-r02h}; mkdir -p $O
NG=$(nvidia-smi -L | wc -l); echo "gpus: $NG"
if [ -n "$2" ]; then timeout 900 python -m pytest tests/test_gpu_comm.py tests/test_gpu_parity.py -x -q 2>&1 | tail -3; fi
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_ref.json 2> $O/${T}_bench_ref.err; echo "ref rc=$?"; head -c 300 $O/${T}_bench_ref.json; echo
for N in 1 2 4 8; do
  [ "$N" -gt "$NG" ] && break
  if [ "$N" -eq 1 ]; then CMD="python bench.py --gpus 1"; else CMD="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N"; fi
  timeout 900 $CMD --steps 10 --warmup 3 > $O/${T}_bench_n$N.json 2> $O/${T}_bench_n$N.err; echo "bench n$N rc=$?"; tail -2 $O/${T}_bench_n$N.err | cut -c1-300
  python - <<PY
import json
try:
    d=json.load(open("$O/${T}_bench_n$N.json"))
    print("N=$N value %.3e e2e %.3e ms %.2f loc %s loc_e2e %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d.get("localizations_per_s"), d.get("localizations_per_s_e2e")))
    print("  e2e breakdown", d["e2e"].get("breakdown_ms"))
    print("  sweep plan", d["roofline"]["splits"], "evals/s/gpu %.3e" % d["roofline"]["evals_per_s"])
    print("  parity", json.dumps(d.get("parity"))[:400])
    print("  comm", d.get("comm", {}).get("nranks"), d.get("comm", {}).get("collectives_issued"))
except Exception as e: print("parse failed", e)
PY
done
