#!/bin/bash
# ncu --set full of one kernel at scale 0.05 (plain run first), source page exported: tools/r2e.sh <tag> <kernel regex>
O=gpurun_out; mkdir -p $O; TAG=$1; K=$2
CMD="python bench.py --scale 0.05 --steps 1 --warmup 1 --no-e2e --no-cpu --no-mask --no-ingest --no-parity"
$CMD > $O/${TAG}_plain.json 2> $O/${TAG}_plain.err && \
ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 -o $O/${TAG}_full -f $CMD > $O/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i $O/${TAG}_full.ncu-rep --page raw --csv > $O/${TAG}_raw.csv 2>/dev/null
ncu -i $O/${TAG}_full.ncu-rep --page source --csv > $O/${TAG}_source.csv 2>/dev/null
rm -f $O/${TAG}_full.ncu-rep
ls -la $O | grep $TAG
