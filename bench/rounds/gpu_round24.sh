#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_engine_gpu.py -m gpu -q --timeout 300 > gpurun_out/tests24.log 2>&1
tail -3 gpurun_out/tests24.log
timeout 120 python bench/ncu_target.py fprop3 fprop5 fprop1 wgrad3 bnbwd3 > gpurun_out/ncu_plain24.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'conv_persist|conv_wgrad_kernel|bn_bwd_fused' -s 10 -c 10 -o gpurun_out/prof24 python bench/ncu_target.py fprop3 fprop5 fprop1 wgrad3 bnbwd3 > gpurun_out/ncu24.log 2>&1
tail -2 gpurun_out/ncu24.log
