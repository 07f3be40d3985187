#!/bin/bash
# One GPU-box session: parity tests, smoke, bench (no profiler), then ncu launch list + one full capture.
# usage: tools/gpu_check.sh <tag> [kernel-regex for the full capture]
TAG=${1:-run}; KREGEX=${2:-k_merge4}
O=gpurun_out; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
tail -3 $O/${TAG}_pytest.log
python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/${TAG}_smoke.log
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; cat $O/${TAG}_bench.json; tail -3 $O/${TAG}_bench.err
if [ -z "$NO_NCU" ]; then
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 600 --csv --log-file $O/${TAG}_launches.csv \
  timeout 900 python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu --no-mask --no-ingest > $O/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s 1 -c 1 -o $O/${TAG}_full -f \
  python bench.py --scale 0.05 --steps 1 --warmup 1 --no-e2e --no-cpu --no-mask --no-ingest > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i $O/${TAG}_full.ncu-rep --page raw --csv > $O/${TAG}_full_raw.csv 2>/dev/null
fi
