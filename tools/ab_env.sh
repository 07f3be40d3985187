# A/B over an environment variable: tools/ab_env.sh VAR v1 v2 ...
VAR=$1; shift
for v in "$@"; do
  env $VAR=$v python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu --no-mask --no-ingest 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$VAR=$v', round(d['ms_per_step'],2), {k:round(x,2) for k,x in d['stage_ms'].items() if isinstance(x,float)}, d['config']['filtered_rows'])"
done
