#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest34.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest34.log
timeout 200 python bench/small_kernels.py > gpurun_out/small34.log 2>&1; cat gpurun_out/small34.log
timeout 300 python bench.py --steps 200 --warmup 20 > gpurun_out/bench34.json 2> gpurun_out/bench34.err; cat gpurun_out/bench34.json
