#!/bin/bash
# A/B of library variants on the default workload: tools/ab_lib.sh name[:ENV=VAL] ...   (name "base" = the in-tree library)
for v in "$@"; do
  n=${v%%:*}; e=${v#*:}; [ "$e" = "$v" ] && e="DK_NOP=1"
  lib=""; [ "$n" != base ] && [ -f build/var/lib_$n.so ] && lib="DK_LIB=build/var/lib_$n.so"
  env $e $lib python bench.py --steps 4 --warmup 3 --no-e2e --no-cpu --no-mask --no-ingest --no-other --no-parity 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); pc=d.get('parity_check',{})
print('$n', round(d['ms_per_step'],2), {k:round(x,2) for k,x in d['stage_ms'].items() if isinstance(x,float)}, 'parity', pc.get('equal'))"
done
