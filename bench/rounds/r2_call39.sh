#!/bin/bash
# round 2, call 39 (1 GPU): same-box A/B of the step: in-kernel split-K sum on/off, BatchNorm tail from TMEM on/off
mkdir -p gpurun_out
for cfg in "1 1" "0 1" "1 0" "0 0" "1 1" "0 1"; do
  set -- $cfg
  DDPB_FUSE_SPLIT=$1 DDPB_TAIL_TMEM=$2 timeout 600 python bench.py --steps 300 --warmup 30 --no-extra > gpurun_out/r2c39_b.json 2> gpurun_out/r2c39_b.err
  python -c "
import json; d=json.load(open('gpurun_out/r2c39_b.json')); print('fuse_split $1 tail_tmem $2:', d['ms_per_step'], d['e2e']['ms_per_step'])" | tee -a gpurun_out/r2c39_ab.txt
done
