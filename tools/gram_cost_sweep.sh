#!/bin/bash
# Developer tool: sweep the planner's relative tile costs (DVF_GRAM_COST / DVF_GRAM_DIAG) for the parts schedule.
big=${1:-"1000000 500 30"}
for full in 240 256 272; do
  for half in 58 64; do
    for s3 in 40 44 48; do
      for diag in 0.65 0.75; do
        echo -n "full=$full half=$half s3=$s3 diag=$diag : "
        DVF_GRAM_NOCHECK=1 DVF_GRAM_PCOST="$full,$half,128,64,$s3,32,16" DVF_GRAM_DIAG=$diag python tools/gram_tune.py $big 2>&1 | tail -1 | cut -c1-60
      done
    done
  done
done
