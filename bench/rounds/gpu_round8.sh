#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q --timeout 120 -k "pair" > gpurun_out/tests8a.log 2>&1
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -q --timeout 300 -k "not pair" > gpurun_out/tests8.log 2>&1
timeout 300 python bench/kernel_bench.py --batch 256 --out gpurun_out/kernel_bench8.json > gpurun_out/kernel_bench8.log 2>&1
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/bench_ours_8.json 2> gpurun_out/bench_ours_8.err
tail -15 gpurun_out/tests8a.log; tail -5 gpurun_out/tests8.log; grep -E "fprop|dgrad|wgrad|totals" gpurun_out/kernel_bench8.log; cut -c1-260 gpurun_out/bench_ours_8.json; tail -3 gpurun_out/bench_ours_8.err
