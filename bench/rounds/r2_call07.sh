#!/bin/bash
# round 2, call 7: prefetch + published loss: tests and end-to-end bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c7_pytest.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2c7_pytest.log | cut -c1-250
for v in "" ""; do
  timeout 200 python bench.py --steps 40 --warmup 10 $v 2> gpurun_out/r2c7_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_last'])"
done
tail -5 gpurun_out/r2c7_bench.err
