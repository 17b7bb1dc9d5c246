#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-graph > gpurun_out/plain6.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 140 --csv --log-file gpurun_out/launches6.csv python bench.py --steps 2 --warmup 3 --no-graph > gpurun_out/ncu6.log 2>&1
tail -2 gpurun_out/ncu6.log; wc -l gpurun_out/launches6.csv
