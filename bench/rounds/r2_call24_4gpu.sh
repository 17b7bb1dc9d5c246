#!/bin/bash
# round 2, call 24 (4 GPUs): new ddp defaults (balanced buckets, 32 x 256 comm CTAs, release-store barrier, deep wgrad ring)
mkdir -p gpurun_out
DDPB_TEST_WORLD=4 timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q -p no:cacheprovider -k "ddp or three_ranks" > gpurun_out/r2c24_pytest.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/r2c24_pytest.log | cut -c1-300
run() { timeout 300 python bench.py --gpus 4 --steps 40 --warmup 10 --no-extra "$@" 2> gpurun_out/r2c24_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['e2e']['ms_per_step'])" || tail -5 gpurun_out/r2c24_bench.err; }
run; run
run --bucket-layout reference --comm-blocks 16 --comm-threads 512
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29611 bench/step_timeline.py > gpurun_out/r2c24_timeline.log 2>&1; grep -v "conv+bn fwd\|^\*\|^W1\|OMP\|^$" gpurun_out/r2c24_timeline.log | tail -48
