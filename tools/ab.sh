#!/bin/bash
# same-box A/B: current tree vs the tree exported under tools/old
for k in 1 2; do
  echo "--- current"; STEPS=8 tools/quick_bench.sh
  echo "--- old"; (cd tools/old && STEPS=8 tools/quick_bench.sh)
done
