#!/bin/bash
# round 2, call 32 (1 GPU): phase timeline of the persistent conv kernel, BatchNorm tail from TMEM vs from L2
mkdir -p gpurun_out
timeout 300 python bench/conv_probe2.py 2,0 > gpurun_out/r2c32_probe_tmem.txt 2>&1; echo "rc=$?"; cat gpurun_out/r2c32_probe_tmem.txt
timeout 300 python bench/conv_probe2.py --tail-l2 2,0 > gpurun_out/r2c32_probe_l2.txt 2>&1; echo "rc=$?"; cat gpurun_out/r2c32_probe_l2.txt
