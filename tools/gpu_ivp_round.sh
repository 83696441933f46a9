#!/bin/bash
# GPU-box visit for K11 (trajectory integration): its parity tests, throughput, one ncu --set full capture, then the
# whole GPU suite and the default bench line.   Usage: bash tools/gpu_ivp_round.sh <tag>
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_ivp.py -x -q -s > $out/pytest_ivp_$tag.log 2>&1; echo "pytest_ivp rc=$?" >> $out/pytest_ivp_$tag.log
tail -5 $out/pytest_ivp_$tag.log
timeout 200 python tools/ivp_tune.py 4096 30 500 250 > $out/ivp_tune_$tag.txt 2>&1
timeout 100 python tools/ivp_tune.py 1024 2 100 250 >> $out/ivp_tune_$tag.txt 2>&1
cat $out/ivp_tune_$tag.txt
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_ivp_rk45 -c 1 -f -o $out/k_ivp_rk45_$tag \
    python tools/ivp_tune.py 2048 30 500 250 > $out/ncu_k_ivp_$tag.log 2>&1
timeout 420 python -m pytest tests -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest rc=$?" >> $out/pytest_$tag.log
tail -3 $out/pytest_$tag.log
timeout 240 python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?" >> $out/bench_$tag.err
cat $out/bench_$tag.json
