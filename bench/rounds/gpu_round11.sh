#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q --timeout 120 -k "variants" > gpurun_out/tests11.log 2>&1
tail -6 gpurun_out/tests11.log
timeout 300 python bench/conv_timeline.py 2,64,1 2,128,1 2,256,1 2,128,2 2,128,4 > gpurun_out/timeline11.log 2>&1
cat gpurun_out/timeline11.log
timeout 300 python bench/conv_probe.py 0,128,1 2,128,1 2,256,1 2,128,2 2,128,4 2,256,2 --epi --warm > gpurun_out/probe11.log 2>&1
cat gpurun_out/probe11.log
