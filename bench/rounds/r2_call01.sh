#!/bin/bash
# round 2, call 1: gated experimental kernels on hardware, per-layer timings of the CTA-pair modes, baseline bench
mkdir -p gpurun_out
export DDPB_TEST_EXPERIMENTAL=1
for t in test_persistent_cta_pair_conv test_persistent_cta_pair_conv_with_batchnorm_tail test_cta_pair_wgrad; do
  timeout 120 python -m pytest tests/test_experimental_gpu.py -m gpu -q -p no:cacheprovider -k "$t and not tail_x" > gpurun_out/r2c1_$t.log 2>&1
  echo "== $t rc=$?"; tail -8 gpurun_out/r2c1_$t.log | cut -c1-220
done
unset DDPB_TEST_EXPERIMENTAL
timeout 200 python bench/layer_graph_times.py > gpurun_out/r2c1_layers_default.log 2>&1; cat gpurun_out/r2c1_layers_default.log
timeout 200 python bench/layer_graph_times.py --conv-mode 3 --block-n 256 > gpurun_out/r2c1_layers_m3_256.log 2>&1; cat gpurun_out/r2c1_layers_m3_256.log
timeout 200 python bench/layer_graph_times.py --conv-mode 3 --block-n 128 > gpurun_out/r2c1_layers_m3_128.log 2>&1; cat gpurun_out/r2c1_layers_m3_128.log
timeout 200 python bench/layer_graph_times.py --wgrad-mode 2 2>&1 | sed 's/| dgrad.*//' > gpurun_out/r2c1_layers_wg2.log; cat gpurun_out/r2c1_layers_wg2.log
timeout 200 python bench.py --steps 20 --warmup 5 > gpurun_out/r2c1_bench.json 2> gpurun_out/r2c1_bench.err; cat gpurun_out/r2c1_bench.json
