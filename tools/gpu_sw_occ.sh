#!/bin/bash
for b in 0 1 2 3 4; do
  if [ $b = 0 ]; then unset GRT_SW_BLOCKS_PER_SM; else export GRT_SW_BLOCKS_PER_SM=$b; fi
  python tools/step_timeline.py 2>&1 | grep "^sw" | grep -o "solver [0-9.]* ms" | sed "s/^/  blocks_per_sm=$b /"
done
