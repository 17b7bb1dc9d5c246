#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -q --timeout 300 > gpurun_out/tests7.log 2>&1
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/bench_ours_7.json 2> gpurun_out/bench_ours_7.err
timeout 300 python bench.py --steps 2 --warmup 3 --no-graph > gpurun_out/plain7.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 140 --csv --log-file gpurun_out/launches7.csv python bench.py --steps 2 --warmup 3 --no-graph > gpurun_out/ncu7.log 2>&1
tail -8 gpurun_out/tests7.log; cut -c1-260 gpurun_out/bench_ours_7.json; tail -3 gpurun_out/bench_ours_7.err
