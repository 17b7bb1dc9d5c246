#!/bin/bash
# round 2, call 41 (1 GPU): %globaltimer stamps between the kernels of the captured step (current build)
mkdir -p gpurun_out
timeout 600 python bench/step_timeline.py > gpurun_out/r2c41_timeline.txt 2>&1; echo "rc=$?"; cat gpurun_out/r2c41_timeline.txt | tail -80
