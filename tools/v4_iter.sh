#!/bin/bash
# one quick GPU iteration on the merge kernel: join parity tests, phase shares (profiling variant), timings at 0.1 and full scale
TAG=${1:-it}; O=gpurun_out; mkdir -p $O
timeout 420 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k join > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
if [ -f build/var/lib_m4prof.so ]; then
DK_LIB=$PWD/build/var/lib_m4prof.so timeout 200 python bench.py --scale 0.1 --steps 2 --warmup 1 --no-e2e --no-cpu --no-mask --no-ingest 2>&1 >/dev/null | grep "m4 prof" | tail -1
fi
for sc in ${SCALES:-0.1 1.0}; do
  timeout 300 python bench.py --scale $sc --steps 3 --warmup 2 --no-e2e --no-cpu --no-mask --no-ingest 2>$O/${TAG}_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('scale', $sc, round(d['ms_per_step'],2), {k:round(v,2) for k,v in d['stage_ms'].items()}, d['config']['filtered_rows'], d['config']['union_kmers'], round(d['roofline']['frac'],4))"
done
