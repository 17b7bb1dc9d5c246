#!/bin/bash
# round 2, call 38 (1 GPU): layer times and step after making the split-K tail mapping a compile-time choice
mkdir -p gpurun_out
timeout 300 python bench/layer_graph_times.py > gpurun_out/r2c38_layers.txt 2>&1; echo "rc=$?"; cat gpurun_out/r2c38_layers.txt
timeout 600 python bench.py --steps 200 --warmup 20 --no-extra > gpurun_out/r2c38_bench.json 2> gpurun_out/r2c38_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c38_bench.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'])"
