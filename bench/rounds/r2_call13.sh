#!/bin/bash
# round 2, call 13: one-epoch parity, with the reference's own spread (TF32 vs fp32 convolutions) and a second run of ours
mkdir -p gpurun_out
timeout 900 python bench/parity_epoch.py --spread --quiet --json gpurun_out/r2c13_parity_a.json 2>&1 | grep -v Warn | tail -2
timeout 900 python bench/parity_epoch.py --spread --ref-fp32 --quiet --json gpurun_out/r2c13_parity_b.json 2>&1 | grep -v Warn | tail -2
