#!/bin/bash
# Developer tool: sweep the planner's relative tile costs (DVF_GRAM_COST / DVF_GRAM_DIAG) for the parts schedule.
big="1000000 500 30"
for half in 64 70.4 78; do
  for s3 in 48 55.2 64 72; do
    for diag in 0.70 0.78 0.86; do
      echo -n "half=$half s3=$s3 diag=$diag : "
      DVF_GRAM_NOCHECK=1 DVF_GRAM_COST="256,$half,134.4,70.4,$s3,40,25.6" DVF_GRAM_DIAG=$diag python tools/gram_tune.py $big 2>&1 | tail -1 | cut -c1-60
    done
  done
done
