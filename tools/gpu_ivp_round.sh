#!/bin/bash
# GPU-box visit for K11 (trajectory integration): parity tests, throughput, then the whole GPU suite, smoke and the
# default bench line.   Usage: bash tools/gpu_ivp_round.sh <tag>   (VARIANTS=1 also times the kernel variants, NCU=1
# captures ncu --set full of the lock-step kernel)
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
timeout 400 python -m pytest tests/test_gpu_ivp.py -x -q -s > $out/pytest_ivp_$tag.log 2>&1; echo "pytest_ivp rc=$?" >> $out/pytest_ivp_$tag.log
tail -8 $out/pytest_ivp_$tag.log
: > $out/ivp_tune_$tag.txt
if [ -n "$VARIANTS" ]; then
for v in 1 2 3; do
  echo "== DVF_IVP_VARIANT=$v" >> $out/ivp_tune_$tag.txt
  DVF_IVP_VARIANT=$v timeout 200 python tools/ivp_tune.py 4096 30 500 250 >> $out/ivp_tune_$tag.txt 2>&1
done
fi
echo "== default" >> $out/ivp_tune_$tag.txt
timeout 200 python tools/ivp_tune.py 4096 30 500 250 >> $out/ivp_tune_$tag.txt 2>&1
timeout 200 python tools/ivp_tune.py 4096 50 1000 100 >> $out/ivp_tune_$tag.txt 2>&1
timeout 100 python tools/ivp_tune.py 1024 2 100 250 >> $out/ivp_tune_$tag.txt 2>&1
if [ -z "$QUICK" ]; then timeout 200 python tools/ivp_tune.py 18944 30 500 250 >> $out/ivp_tune_$tag.txt 2>&1; fi
cat $out/ivp_tune_$tag.txt
if [ -n "$NCU" ]; then
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_ivp_rk45_t -c 1 -f -o $out/k_ivp_rk45_t_$tag \
    python tools/ivp_tune.py 4096 30 500 250 > $out/ncu_k_ivp_$tag.log 2>&1
fi
timeout 420 python -m pytest tests -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest rc=$?" >> $out/pytest_$tag.log
tail -3 $out/pytest_$tag.log
timeout 300 python __graft_entry__.py smoke > $out/smoke_$tag.log 2>&1; echo "smoke rc=$?" >> $out/smoke_$tag.log
tail -2 $out/smoke_$tag.log
if [ -z "$QUICK" ]; then
timeout 240 python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?" >> $out/bench_$tag.err
cat $out/bench_$tag.json
fi
