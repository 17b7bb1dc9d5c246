#!/bin/bash
# round 2, call 2: phase timeline + producer-skip probes of the persistent conv kernels
mkdir -p gpurun_out
timeout 300 python bench/conv_probe2.py --probe 2,0 3,256 3,128 > gpurun_out/r2c2_probe_fwd.log 2>&1; cat gpurun_out/r2c2_probe_fwd.log
timeout 300 python bench/conv_probe2.py --dgrad 2,0 3,256 3,128 > gpurun_out/r2c2_probe_dgrad.log 2>&1; cat gpurun_out/r2c2_probe_dgrad.log
