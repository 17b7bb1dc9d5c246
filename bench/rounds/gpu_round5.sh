#!/bin/bash
mkdir -p gpurun_out
timeout 200 python bench/diag_layers.py > gpurun_out/diag_layers.log 2>&1
timeout 120 python bench/ncu_target.py fprop3 fprop5 wgrad3 bnbwd0 > gpurun_out/ncu_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'conv_fprop|conv_wgrad_kernel|bn_bwd' -c 15 -o gpurun_out/prof5 python bench/ncu_target.py fprop3 fprop5 wgrad3 bnbwd0 > gpurun_out/ncu5.log 2>&1
cat gpurun_out/diag_layers.log; tail -3 gpurun_out/ncu5.log
