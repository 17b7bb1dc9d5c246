#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest39.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest39.log
timeout 300 python bench.py --steps 200 --warmup 20 > gpurun_out/bench39.json 2> gpurun_out/bench39.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench39.json; tail -3 gpurun_out/bench39.err
timeout 300 python bench/layer_graph_times.py > gpurun_out/layers39.log 2>&1; cat gpurun_out/layers39.log
