#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/x_build.log 2>&1
for s in "32768 320 320" "2048 1280 1280" "8192 640 640"; do
  set -- $s
  echo "=== M=$1 K=$2 N=$3 (cta 4, warm) ==="
  timeout 120 python tools/timeline.py --M $1 --K $2 --N $3 --cta 4 2>&1 | tail -40
done > gpurun_out/x_timeline.txt 2>&1
cat gpurun_out/x_timeline.txt
