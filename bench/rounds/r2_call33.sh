#!/bin/bash
# round 2, call 33 (1 GPU): BatchNorm tail of the persistent conv kernel from TMEM vs from L2 (numerics, layer times, step)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -p no:cacheprovider  > gpurun_out/r2c33_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2c33_pytest.log | cut -c1-300
timeout 300 python bench/layer_graph_times.py > gpurun_out/r2c33_layers_tmem.txt 2>&1; echo "rc=$?"; cat gpurun_out/r2c33_layers_tmem.txt | cut -c1-60
timeout 300 python bench/layer_graph_times.py --tail-l2 > gpurun_out/r2c33_layers_l2.txt 2>&1; echo "rc=$?"; cat gpurun_out/r2c33_layers_l2.txt | cut -c1-60
timeout 600 python bench.py --steps 200 --warmup 20 --no-extra > gpurun_out/r2c33_bench.json 2> gpurun_out/r2c33_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c33_bench.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'])"
