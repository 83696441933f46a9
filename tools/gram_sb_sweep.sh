#!/bin/bash
# Developer tool: block length / lag of the parts schedule.
big=${1:-"1000000 500 30"}
for sbp in 256 512 1024 2048; do for lag in 0 1; do
  echo -n "sbp=$sbp lag=$lag : "; DVF_GRAM_NOCHECK=1 DVF_GRAM_SBP=$sbp DVF_GRAM_LAG=$lag python tools/gram_tune.py $big 2>&1 | tail -1 | cut -c1-60
done; done
