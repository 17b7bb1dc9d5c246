#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -q --timeout 300 > gpurun_out/tests4.log 2>&1
timeout 300 python bench/kernel_bench.py --batch 256 --out gpurun_out/kernel_bench4.json > gpurun_out/kernel_bench4.log 2>&1
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/bench_ours_4.json 2> gpurun_out/bench_ours_4.err
tail -12 gpurun_out/tests4.log; cat gpurun_out/kernel_bench4.log; cut -c1-300 gpurun_out/bench_ours_4.json; tail -3 gpurun_out/bench_ours_4.err
