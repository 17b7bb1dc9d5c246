#!/bin/bash
# round 2, call 10 (2 GPUs): both bench arms with the extra_configs block
mkdir -p gpurun_out
timeout 600 python bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2c10_ours2.json 2> gpurun_out/r2c10_ours2.err; echo "rc=$?"; cat gpurun_out/r2c10_ours2.json | cut -c1-3000; tail -3 gpurun_out/r2c10_ours2.err
timeout 600 python bench.py --gpus 2 --steps 20 --warmup 5 --impl reference > gpurun_out/r2c10_ref2.json 2> gpurun_out/r2c10_ref2.err; echo "rc=$?"; cat gpurun_out/r2c10_ref2.json | cut -c1-3000; tail -3 gpurun_out/r2c10_ref2.err
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2c10_ours1.json 2> gpurun_out/r2c10_ours1.err; echo "rc=$?"; cat gpurun_out/r2c10_ours1.json | cut -c1-600
