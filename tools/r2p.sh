#!/bin/bash
# GPU session: evidence for profiles/ — ncu --set full of the two top kernels (scale 0.05, plain run first), the launch
# list of a full-size step, and one full-scale --set full launch of the merge kernel for the measured DRAM traffic.
# usage: tools/r2p.sh <tag>
O=gpurun_out; mkdir -p $O; TAG=$1
SM="python bench.py --scale 0.05 --steps 1 --warmup 1 --no-e2e --no-cpu --no-mask --no-ingest --no-parity"
FS="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-mask --no-ingest --no-parity"
cap() { # name, kernel regex, command...
  n=$1; k=$2; shift 2
  "$@" > $O/${TAG}_${n}_plain.json 2> $O/${TAG}_${n}_plain.err && \
  ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o $O/${TAG}_${n} -f "$@" > $O/${TAG}_${n}_ncu.log 2>&1; echo "ncu $n rc=$?"
  ncu -i $O/${TAG}_${n}.ncu-rep --page raw --csv > $O/${TAG}_${n}_raw.csv 2>/dev/null
  ncu -i $O/${TAG}_${n}.ncu-rep --page source --csv > $O/${TAG}_${n}_source.csv 2>/dev/null
  rm -f $O/${TAG}_${n}.ncu-rep
}
cap merge k_merge5 $SM
cap ttest k_ttest_rows8 $SM
cap mergefull k_merge5 $FS
$FS > $O/${TAG}_fs_plain.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv --log-file $O/${TAG}_launches.csv $FS > $O/${TAG}_launches.log 2>&1; echo "launch list rc=$?"
ls -la $O | grep $TAG | awk '{print $5, $9}'
