#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench (both arms), ncu launch list, ncu --set full captures of the hot kernels.
# Usage (from the repo root, under gpurun): bash tools/gpu_round.sh <tag>
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $out/gpu_$tag.txt
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest rc=$?" >> $out/pytest_$tag.log
timeout 300 python __graft_entry__.py smoke > $out/smoke_$tag.log 2>&1; echo "smoke rc=$?" >> $out/smoke_$tag.log
timeout 900 python bench.py > $out/bench_$tag.json 2> $out/bench_$tag.err; echo "bench rc=$?" >> $out/bench_$tag.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $out/bench_ref_$tag.json 2> $out/bench_ref_$tag.err
if [ -z "$NO_NCU" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $out/launches_$tag.csv \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $out/ncu_list_$tag.log 2>&1
for k in k_gram_parts k_gram k_jacobian_gemm k_dac k_update_v k_chol_solve; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^$k\$" --launch-skip 1 -c 1 -f -o $out/${k}_$tag \
      python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > $out/ncu_${k}_$tag.log 2>&1
done
fi
tail -3 $out/pytest_$tag.log; tail -2 $out/smoke_$tag.log; cat $out/bench_$tag.json; cat $out/bench_ref_$tag.json
