#!/bin/bash
# round 2, call 5: deterministic two-kernel BatchNorm backward: tests, per-layer times, A/B bench, timeline
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_engine_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c5_pytest.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/r2c5_pytest.log | cut -c1-250
for m in 3 2; do timeout 200 python bench/layer_graph_times.py --bn-fused $m 2>&1 | sed 's/| wgrad.*//' > gpurun_out/r2c5_layers_bn$m.log; cat gpurun_out/r2c5_layers_bn$m.log; done
for v in "" "--bn-fused 2" "--no-overlap-sgd" "--bn-fused 2 --no-overlap-sgd"; do
  timeout 200 python bench.py --steps 40 --warmup 10 $v 2> gpurun_out/r2c5_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_last'])"
done
timeout 200 python bench/step_timeline.py > gpurun_out/r2c5_timeline.log 2>&1; cat gpurun_out/r2c5_timeline.log
