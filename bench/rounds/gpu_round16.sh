#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py tests/test_zz_no_alloc.py -m gpu -q --timeout 300 > gpurun_out/tests16.log 2>&1
tail -4 gpurun_out/tests16.log
timeout 300 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours_16.json 2> gpurun_out/bench_ours_16.err
cut -c1-260 gpurun_out/bench_ours_16.json; tail -3 gpurun_out/bench_ours_16.err
