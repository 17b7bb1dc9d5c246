#!/bin/bash
# round 2, call 28 (1 GPU): full GPU suite with the device-side augmentation RNG, (tile width, split-K) sweep against the planner, bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2c28_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2c28_pytest.log | cut -c1-300
timeout 600 python bench/conv_plan_sweep.py > gpurun_out/r2c28_plan_sweep.txt 2>&1; echo "sweep rc=$?"; grep -E "^L|possible" gpurun_out/r2c28_plan_sweep.txt
timeout 600 python bench.py --steps 200 --warmup 20 --no-extra > gpurun_out/r2c28_bench.json 2> gpurun_out/r2c28_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c28_bench.json')); print(d['ms_per_step'], d['e2e'], d['clocks'])"
