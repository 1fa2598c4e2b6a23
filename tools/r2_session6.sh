#!/bin/bash
O=gpurun_out; T=${1:-r02f}; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${T}_pytest.log
timeout 300 python bench.py --precision f64 --steps 3 --no-localization --no-extra > $O/${T}_bench_f64.json 2> $O/${T}_bench_f64.err; echo "bench f64 rc=$?"; tail -2 $O/${T}_bench_f64.err; head -c 400 $O/${T}_bench_f64.json; echo
timeout 900 python bench.py > $O/${T}_bench_n1.json 2> $O/${T}_bench_n1.err; echo "bench n1 rc=$?"; tail -3 $O/${T}_bench_n1.err
python - <<PY
import json
d=json.load(open("$O/${T}_bench_n1.json"))
print("value",d["value"],"e2e",d["e2e"]["value"],"loc",d.get("localizations_per_s"),d.get("localizations_per_s_e2e"))
print("parity",json.dumps(d.get("parity"))[:900])
print("f64",json.dumps(d.get("f64"))[:700])
print("other",json.dumps(d.get("other_sweeps"))[:900])
print("table",json.dumps(d.get("configs_table"))[:1500])
print("inlier",d["roofline"]["inlier_pass"])
PY
