#!/bin/bash
# K7 with parked kernel values against the recomputing version: parity tests, then the stage timing of the bench.
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
timeout 300 python -m pytest tests -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest rc=$?" >> $out/pytest_$tag.log
tail -3 $out/pytest_$tag.log
timeout 120 python bench.py --no-e2e --no-cpu-baseline > $out/bench_park_$tag.json 2> $out/bench_park_$tag.err
DVF_DAC_NOPARK=1 timeout 120 python bench.py --no-e2e --no-cpu-baseline > $out/bench_nopark_$tag.json 2> $out/bench_nopark_$tag.err
for k in park nopark; do
  python -c "import json,sys; j=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], j['value'], j['ms_per_step'], j['stages_ms'])" $out/bench_${k}_$tag.json $k
done
