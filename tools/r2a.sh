#!/bin/bash
# GPU session r2a: v5 merge kernel — parity first, then A/B timings of its variants against v4
O=gpurun_out; mkdir -p $O
timeout 120 python __graft_entry__.py smoke > $O/r2a_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/r2a_smoke.log
timeout 900 python -m pytest tests -m gpu -q > $O/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $O/r2a_pytest.log
B="python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-mask --no-ingest"
run() { # name, env...
  n=$1; shift
  env "$@" timeout 400 $B > $O/r2a_bench_$n.json 2> $O/r2a_bench_$n.err; echo "bench $n rc=$?"
  python - "$O/r2a_bench_$n.json" <<'PY'
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
  grep -h "m5 prof\|m4 prof" $O/r2a_bench_$n.err | tail -1
}
run v5
run v4 DK_MERGE=4
run v5prof DK_LIB=build/var/lib_m5prof.so
run v5mb8 DK_LIB=build/var/lib_m5mb8.so
run v5vt9 DK_LIB=build/var/lib_m5vt9.so
run v5vt7 DK_LIB=build/var/lib_m5vt7.so
run v5ctas2 DK_M5_CTAS=2
B="python bench.py --config 5 --steps 3 --warmup 2 --no-e2e --no-cpu --no-mask --no-ingest"
run c5v5
run c5v3 DK_MERGE=3
run c5v4 DK_MERGE=4
tail -5 $O/r2a_bench_v5.err
