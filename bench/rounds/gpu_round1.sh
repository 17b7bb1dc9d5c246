#!/bin/bash
# First GPU contact: kernel numerics per group (separate processes so one trap does not
# poison the rest), then per-layer timings.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
for grp in fprop dgrad wgrad bn_relu fc_cross sgd pack first_layer_operand; do
  echo "=== $grp" >> gpurun_out/tests1.log
  timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "$grp" --timeout 120 -x 2>&1 | tail -40 >> gpurun_out/tests1.log
done
timeout 300 python bench/kernel_bench.py --batch 256 --out gpurun_out/kernel_bench.json > gpurun_out/kernel_bench.log 2>&1
tail -5 gpurun_out/tests1.log
tail -30 gpurun_out/kernel_bench.log
