#!/bin/bash
# ncu visit: launch list of one bench step + --set full captures of the named kernels.  Usage: bash tools/ncu_round.sh <tag> "<kernels>" [bench args]
tag=${1:-r02}; kernels=${2:-"k_gram_parts k_lu_solve k_dac k_update_v k_kernel_matrix"}; shift 2
out=gpurun_out
args="--steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-parity --no-variants $@"
python bench.py $args > $out/bench_plain_$tag.json 2> $out/bench_plain_$tag.err || { echo "plain bench failed"; tail -5 $out/bench_plain_$tag.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $out/launches_$tag.csv \
    python bench.py $args > $out/ncu_list_$tag.log 2>&1
for k in $kernels; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"^$k" --launch-skip 1 -c 1 -f -o $out/${k}_$tag \
      python bench.py $args > $out/ncu_${k}_$tag.log 2>&1
done
ls -la $out/*_$tag.ncu-rep
