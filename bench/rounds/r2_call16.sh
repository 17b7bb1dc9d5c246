#!/bin/bash
# round 2, call 16: channel-owner kernels (BatchNorm backward, split-K finalize + BatchNorm forward), deeper BN tail,
# after the removal of the losing kernel variants: full GPU suite, layer times, bench, timeline
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c16_pytest.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r2c16_pytest.log | cut -c1-250
timeout 200 python bench/layer_graph_times.py > gpurun_out/r2c16_layers.log 2>&1; cat gpurun_out/r2c16_layers.log
for v in "" ""; do
  timeout 200 python bench.py --steps 40 --warmup 10 $v 2> gpurun_out/r2c16_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_last'], d['gpu_launches_per_step'])"
done
