#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py tests/test_zz_no_alloc.py -m gpu -q --timeout 300 > gpurun_out/tests22.log 2>&1
tail -5 gpurun_out/tests22.log
timeout 300 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours_22.json 2> gpurun_out/bench_ours_22.err
cut -c1-260 gpurun_out/bench_ours_22.json; tail -3 gpurun_out/bench_ours_22.err
timeout 300 python bench.py --steps 2 --warmup 3 --no-graph > gpurun_out/plain22.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 160 --csv --log-file gpurun_out/launches22.csv python bench.py --steps 2 --warmup 3 --no-graph > gpurun_out/ncu22.log 2>&1
