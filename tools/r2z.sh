#!/bin/bash
# final evidence of a round on one B200: GPU suite, smoke, default bench, reference arm (then tools/r2p.sh for the ncu captures)
O=gpurun_out; mkdir -p $O; TAG=$1
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/${TAG}_smoke.log
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/${TAG}_pytest.log
python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_reference.json 2> $O/${TAG}_reference.err; echo "reference rc=$?"
python - "$O/${TAG}_bench.json" "$O/${TAG}_reference.json" <<'PY'
import sys, json
d = json.loads(open(sys.argv[1]).read()); r = json.loads(open(sys.argv[2]).read())
print("value %.4e ms %.2f e2e %.4e frac %.3f parity %s launches %d" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["parity_check"]["equal"], d["gpu_launches"]))
print("stages", {k: round(v, 2) for k, v in d["stage_ms"].items() if isinstance(v, float)})
print("cpu1", d["cpu_baseline"]["value"], d["cpu_baseline"]["stage_seconds"], "port", d["cpu_baseline_port"]["value"], d["cpu_baseline_port"]["agrees_with_reference_tools"])
print("cfg5", d["other_configs"]["cfg5"]["value"], d["other_configs"]["cfg5"]["ms_per_step"], d["other_configs"]["cfg5"]["full_size_invariants_equal"])
print("ref", r["value"], r["ms_per_step"], r["cpu_baseline"]["cores"], r["cpu_baseline"]["stage_seconds_last_step"])
PY
