#!/bin/bash
# round 2, call 8 (2 GPUs): multi-GPU tests with the race detector, ddp A/B (PDL, overlap SGD)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c8_pytest.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2c8_pytest.log | cut -c1-300
run() { timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 "$@" 2> gpurun_out/r2c8_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_last'])" || tail -5 gpurun_out/r2c8_bench.err; }
run
run --no-overlap-sgd
DDPB_PDL=1 run
DDPB_PDL=1 run --no-overlap-sgd
run --mode allreduce
run --mode gather_scatter
run --comm-blocks 32
