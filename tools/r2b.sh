#!/bin/bash
# GPU session: merge-kernel iteration — join parity tests, then A/B timings (tag = $1, variants = rest)
TAG=$1; shift
O=gpurun_out; mkdir -p $O
timeout 120 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/${TAG}_smoke.log
timeout 900 python -m pytest tests -m gpu -q -k "join or config1 or sharded or ingest" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-mask --no-ingest"
run() { # name, env...
  n=$1; shift
  env "$@" timeout 400 $B > $O/${TAG}_bench_$n.json 2> $O/${TAG}_bench_$n.err; echo "bench $n rc=$?"
  python - "$O/${TAG}_bench_$n.json" <<'PY'
import sys, json
try:
    d = json.loads(open(sys.argv[1]).read())
    pc = d.get("parity_check", {})
    print("   ms/step %.2f  merge %.2f  stages %s  frac %.3f  parity %s/%s" % (d["ms_per_step"], d["stage_ms"]["merge_kernel"],
          {k: round(v, 2) for k, v in d["stage_ms"].items() if isinstance(v, float)}, d["roofline"]["frac"],
          pc.get("full_size", {}).get("equal"), pc.get("sample_equal")))
except Exception as e:
    print("   no result:", e)
PY
  grep -h "m5 prof\|m4 prof" $O/${TAG}_bench_$n.err | tail -1
}
for v in "$@"; do
  n=${v%%:*}; e=${v#*:}
  if [ "$e" = "$v" ] || [ -z "$e" ]; then run $n DK_NOP=1; else run $n $e; fi
done
