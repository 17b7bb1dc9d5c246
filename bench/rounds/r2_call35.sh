#!/bin/bash
# round 2, call 35 (1 GPU): phase timeline of the split-K layers with the in-kernel sum
mkdir -p gpurun_out
timeout 300 python bench/conv_probe2.py 2,0 > gpurun_out/r2c35_probe_fwd.txt 2>&1; echo "rc=$?"; grep -E "^L[4-7]" gpurun_out/r2c35_probe_fwd.txt
timeout 300 python bench/conv_probe2.py --dgrad 2,0 > gpurun_out/r2c35_probe_dgrad.txt 2>&1; echo "rc=$?"; grep -E "^L[4-7]" gpurun_out/r2c35_probe_dgrad.txt
