#!/bin/bash
# GPU session: ONE ncu capture. usage: tools/r2t.sh <tag> ttest|launches
O=gpurun_out; mkdir -p $O; TAG=$1
SM="python bench.py --scale 0.05 --steps 1 --warmup 1 --no-e2e --no-cpu --no-mask --no-ingest --no-parity --no-other"
FS="python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-mask --no-ingest --no-parity --no-other"
if [ "$2" = ttest ]; then
  $SM > $O/${TAG}_ttest_plain.json 2> $O/${TAG}_ttest_plain.err && \
  ncu --set full --clock-control none --import-source on -k regex:k_ttest_rows8 -s 1 -c 1 -o $O/${TAG}_ttest -f $SM > $O/${TAG}_ttest_ncu.log 2>&1; echo "ncu ttest rc=$?"
  ncu -i $O/${TAG}_ttest.ncu-rep --page raw --csv > $O/${TAG}_ttest_raw.csv 2>/dev/null
  ncu -i $O/${TAG}_ttest.ncu-rep --page source --csv > $O/${TAG}_ttest_source.csv 2>/dev/null
  rm -f $O/${TAG}_ttest.ncu-rep
else
  $FS > $O/${TAG}_fs_plain.json 2>/dev/null && \
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv --log-file $O/${TAG}_launches.csv $FS > $O/${TAG}_launches.log 2>&1; echo "launch list rc=$?"
fi
ls -la $O | grep $TAG | awk '{print $5, $9}'
