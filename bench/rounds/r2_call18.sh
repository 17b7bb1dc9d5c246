#!/bin/bash
# round 2, call 18: BatchNorm backward (two-kernel path) at 2 / 3 / 4 CTAs per SM
mkdir -p gpurun_out
for v in "" _bn3 _bn4; do
  echo "== lib$v"
  DDPB_LIB=$PWD/ddpb200/_lib/libddpb200$v.so timeout 200 python bench/layer_graph_times.py 2>&1 | sed 's/| wgrad.*//' | tee gpurun_out/r2c18_layers$v.log | grep -v "^batch"
done
