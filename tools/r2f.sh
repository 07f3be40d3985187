#!/bin/bash
# GPU session: t-test kernel iteration — parity tests touching the t-test, then A/B timings
TAG=$1; shift
O=gpurun_out; mkdir -p $O
timeout 120 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/${TAG}_smoke.log
timeout 900 python -m pytest tests -m gpu -q -k "not join and not ingest and not tags and not mask" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/${TAG}_pytest.log
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-mask --no-ingest"
for v in "$@"; do
  n=${v%%:*}; e=${v#*:}; [ "$e" = "$v" ] && e="DK_NOP=1"; [ -z "$e" ] && e="DK_NOP=1"
  env $e timeout 400 $B > $O/${TAG}_bench_$n.json 2> $O/${TAG}_bench_$n.err; echo "bench $n rc=$?"
  python - "$O/${TAG}_bench_$n.json" <<'PY'
import sys, json
try:
    d = json.loads(open(sys.argv[1]).read())
    pc = d.get("parity_check", {})
    print("   ms/step %.2f  stages %s  parity %s/%s D=%s nferr=%s adjerr=%s" % (d["ms_per_step"],
          {k: round(v, 2) for k, v in d["stage_ms"].items() if isinstance(v, float)},
          pc.get("full_size", {}).get("equal"), pc.get("sample_equal"), pc.get("gpu", {}).get("D"), pc.get("nf_max_rel_err"), pc.get("surv_adj_sum_rel_err")))
except Exception as e:
    print("   no result:", e)
PY
done
