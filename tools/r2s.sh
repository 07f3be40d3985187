#!/bin/bash
# multi-GPU session: tools/r2s.sh <tag> <N> <configs...>   (run under gpurun --gpus N)
TAG=$1; N=$2; shift 2
O=gpurun_out; mkdir -p $O
for C in "$@"; do
  if [ "$N" = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"; fi
  timeout 900 $L bench.py --gpus $N --config $C --steps 3 --warmup 3 --no-cpu --no-mask --no-ingest --no-e2e > $O/${TAG}_cfg${C}_n${N}.json 2> $O/${TAG}_cfg${C}_n${N}.err
  echo "cfg $C N=$N rc=$?"
  python - "$O/${TAG}_cfg${C}_n${N}.json" <<'PY'
import sys, json
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    pc = d.get("parity_check", {})
    print("   value %.3e  ms/step %.2f  stages %s  frac %.3f kernel %s" % (d["value"], d["ms_per_step"],
          {k: round(v, 2) for k, v in d["stage_ms"].items() if isinstance(v, float)}, d["roofline"]["frac"], d["roofline"]["kernel"][:10]))
    print("   U %d R %d D %d; parity full %s sample %s (D gpu %s cpu %s)" % (d["config"]["union_kmers"], d["config"]["filtered_rows"], d["config"]["survivors"],
          pc.get("full_size", {}).get("equal"), pc.get("sample_equal"), pc.get("gpu", {}).get("D"), pc.get("cpu", {}).get("D")))
except Exception as e:
    print("   no result:", e)
PY
  tail -3 $O/${TAG}_cfg${C}_n${N}.err
done
