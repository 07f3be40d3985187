#!/bin/bash
# tools/ab_keys.sh <lib variants...>: default keys and bursts of 8 for each library variant ("base" = in-tree)
for n in "$@"; do
  lib=""; [ "$n" != base ] && lib="DK_LIB=build/var/lib_$n.so"
  for a in "" "--key-bursts 8" "--config 5"; do
    env DK_NOP=1 $lib python bench.py $a --steps 3 --warmup 3 --no-e2e --no-cpu --no-mask --no-ingest --no-other 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); pc=d.get('parity_check',{})
print('$n [$a]', round(d['ms_per_step'],2), 'merge', round(d['stage_ms']['merge_kernel'],2), 'parity', pc.get('equal'))"
  done
done
