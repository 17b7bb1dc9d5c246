#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest35.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest35.log
timeout 300 python bench.py --steps 200 --warmup 20 > gpurun_out/bench35.json 2> gpurun_out/bench35.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench35.json; tail -3 gpurun_out/bench35.err
timeout 300 python bench.py --steps 200 --warmup 20 --no-pdl > gpurun_out/bench35_nopdl.json 2> gpurun_out/bench35_nopdl.err; cut -c1-330 gpurun_out/bench35_nopdl.json
