#!/bin/bash
# round 2, call 4: per-group SGD overlapped with backward (single GPU): tests, A/B bench, timeline
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c4_pytest.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r2c4_pytest.log | cut -c1-250
for v in "" "--no-overlap-sgd" "" "--no-overlap-sgd"; do
  timeout 200 python bench.py --steps 40 --warmup 10 $v 2> gpurun_out/r2c4_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_last'])"
done
timeout 200 python bench/step_timeline.py > gpurun_out/r2c4_timeline.log 2>&1; cat gpurun_out/r2c4_timeline.log
