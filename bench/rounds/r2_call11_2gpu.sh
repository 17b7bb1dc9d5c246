#!/bin/bash
# round 2, call 11 (2 GPUs): batched per-parameter all-reduce + one-shot tail bucket + NVLS root gather
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c11_pytest.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/r2c11_pytest.log | cut -c1-300
run() { timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 --no-extra "$@" 2> gpurun_out/r2c11_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_last'])" || tail -5 gpurun_out/r2c11_bench.err; }
run
run --mode allreduce
run --mode allreduce --granularity param
run --mode allreduce --granularity param_launches
run --mode gather_scatter
run --mode gather_scatter --no-multicast
run --scaling strong --mode allreduce --granularity param
