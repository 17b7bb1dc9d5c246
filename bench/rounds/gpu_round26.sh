#!/bin/bash
mkdir -p gpurun_out
for f in "" "--nostats" "--nostats --nobias"; do
echo "=== flags: $f"
timeout 300 python bench/conv_timeline.py 2,64,1 2,128,1 $f 2>&1 | grep -E "L0|L1"
done > gpurun_out/timeline26.log 2>&1
cat gpurun_out/timeline26.log
