#!/bin/bash
# round 2, call 43 (1 GPU): two-kernel split-K partial tiles through TMA stores: tests, layer times, step
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py tests/test_engine_gpu.py -m gpu -q -x -p no:cacheprovider > gpurun_out/r2c43_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2c43_pytest.log | cut -c1-300
timeout 300 python bench/layer_graph_times.py > gpurun_out/r2c43_layers.txt 2>&1; echo "rc=$?"; cat gpurun_out/r2c43_layers.txt
timeout 600 python bench.py --steps 300 --warmup 30 --no-extra > gpurun_out/r2c43_bench.json 2> gpurun_out/r2c43_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c43_bench.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'])"
