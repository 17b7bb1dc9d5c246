#!/bin/bash
mkdir -p gpurun_out
timeout 200 python bench/small_kernels.py > gpurun_out/small33.log 2>&1; cat gpurun_out/small33.log
