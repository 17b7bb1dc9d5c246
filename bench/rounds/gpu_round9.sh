#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/probe9.log
for v in "" var_S2 var_S6 var_SKA var_SKB; do
  if [ -n "$v" ]; then export DDPB_LIB=$PWD/ddpb200/_lib/$v.so; else unset DDPB_LIB; fi
  timeout 200 python bench/conv_probe.py 0,128,1 >> gpurun_out/probe9.log 2>&1
done
unset DDPB_LIB
timeout 200 python bench/conv_probe.py 1,128,1 1,256,1 0,256,1 0,128,2 1,128,2 1,256,2 >> gpurun_out/probe9.log 2>&1
cat gpurun_out/probe9.log
