#!/bin/bash
# A/B of library variants on one configuration: tools/ab_cfg.sh <config> <extra bench flags or ""> name ...
C=$1; X=$2; shift 2
for n in "$@"; do
  lib=""; [ "$n" != base ] && [ -f build/var/lib_$n.so ] && lib="DK_LIB=build/var/lib_$n.so"
  env DK_NOP=1 $lib python bench.py --config $C --steps 3 --warmup 3 --no-e2e --no-cpu --no-mask --no-ingest --no-other $X 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); pc=d.get('parity_check',{})
print('cfg$C $n', round(d['ms_per_step'],2), {k:round(x,2) for k,x in d.get('stage_ms',{}).items() if isinstance(x,float)}, 'parity', pc.get('equal'))"
done
