#!/bin/bash
# round 2, call 26: BatchNorm backward with the packed-max pooling window: tests, layer times, bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c26_pytest.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r2c26_pytest.log | cut -c1-250
timeout 200 python bench/layer_graph_times.py > gpurun_out/r2c26_layers.log 2>&1; cat gpurun_out/r2c26_layers.log
for v in "" ""; do
  timeout 200 python bench.py --steps 40 --warmup 10 $v 2> gpurun_out/r2c26_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_last'], d['gpu_launches_per_step'])"
done
