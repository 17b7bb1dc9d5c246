#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/conv_timeline.py 0,64,1 0,128,1 0,256,1 1,128,1 1,256,1 0,128,4 > gpurun_out/timeline10.log 2>&1
echo "--- cold" >> gpurun_out/timeline10.log
timeout 300 python bench/conv_timeline.py 0,128,1 1,256,1 --cold >> gpurun_out/timeline10.log 2>&1
cat gpurun_out/timeline10.log
