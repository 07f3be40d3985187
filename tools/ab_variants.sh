for n in pf1_t256 pf1_t512 pf1_t512_tab64 pf1_t1024_tab64 pf1_t1024_tab128 pf2_t512_tab64; do
  DK_LIB=build/var/lib_$n.so python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$n', round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items() if isinstance(v,float)})"
done
