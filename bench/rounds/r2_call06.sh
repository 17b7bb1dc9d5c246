#!/bin/bash
# round 2, call 6: tuned two-kernel BatchNorm backward vs the cooperative kernel
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k bn_relu_pool > gpurun_out/r2c6_pytest.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/r2c6_pytest.log | cut -c1-250
for m in 3 2; do timeout 200 python bench/layer_graph_times.py --bn-fused $m 2>&1 | sed 's/| wgrad.*//' > gpurun_out/r2c6_layers_bn$m.log; cat gpurun_out/r2c6_layers_bn$m.log; done
for v in "" "--bn-fused 2" "" "--bn-fused 2"; do
  timeout 200 python bench.py --steps 40 --warmup 10 $v 2> gpurun_out/r2c6_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_last'])"
done
