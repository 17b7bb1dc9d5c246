#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -q --timeout 300 > gpurun_out/tests13.log 2>&1
tail -5 gpurun_out/tests13.log
timeout 300 python bench/conv_timeline.py 2,0,1 > gpurun_out/timeline13.log 2>&1; cat gpurun_out/timeline13.log
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/bench_ours_13.json 2> gpurun_out/bench_ours_13.err
cut -c1-260 gpurun_out/bench_ours_13.json; tail -3 gpurun_out/bench_ours_13.err
timeout 300 python bench.py --steps 2 --warmup 3 --no-graph > gpurun_out/plain13.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 160 --csv --log-file gpurun_out/launches13.csv python bench.py --steps 2 --warmup 3 --no-graph > gpurun_out/ncu13.log 2>&1
