#!/bin/bash
# round 2, call 37 (1 GPU): split-K summed inside the persistent conv kernel (numerics, plan sweep, layer times, step)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider > gpurun_out/r2c37_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2c37_pytest.log | cut -c1-400
timeout 600 python bench/conv_plan_sweep.py --dgrad-two-kernels > gpurun_out/r2c37_plan_sweep.txt 2>&1; echo "sweep rc=$?"; grep -E "^L|possible" gpurun_out/r2c37_plan_sweep.txt
timeout 300 python bench/layer_graph_times.py > gpurun_out/r2c37_layers.txt 2>&1; echo "rc=$?"; cat gpurun_out/r2c37_layers.txt
timeout 600 python bench.py --steps 200 --warmup 20 --no-extra > gpurun_out/r2c37_bench.json 2> gpurun_out/r2c37_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c37_bench.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'])"
