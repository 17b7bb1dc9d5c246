#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/conv_timeline.py 2,64,1 2,128,1 2,256,1 2>&1 | grep -E "L0|L1|L3|L5" > gpurun_out/timeline31.log
cat gpurun_out/timeline31.log
