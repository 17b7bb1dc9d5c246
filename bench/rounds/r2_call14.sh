#!/bin/bash
# round 2, call 14: distribution of the one-epoch result over 6 seeds per arm
mkdir -p gpurun_out
timeout 1200 python bench/parity_epoch.py --repeat 6 --json gpurun_out/r2c14_parity_dist.json 2>&1 | grep -v Warn | tail -2
