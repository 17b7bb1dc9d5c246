#!/bin/bash
# round 2, call 12: one-epoch parity run, both arms
mkdir -p gpurun_out
timeout 900 python bench/parity_epoch.py --json gpurun_out/r2c12_parity.json > gpurun_out/r2c12_parity.log 2>&1; echo "rc=$?"; grep -v Warning gpurun_out/r2c12_parity.log | tail -40
