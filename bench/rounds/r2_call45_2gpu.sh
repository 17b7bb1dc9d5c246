#!/bin/bash
# round 2, call 45 (2 GPUs): multi-GPU tests and ddp bench with today's kernels (TMEM tail, in-kernel split-K, TMA-store epilogues)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q -p no:cacheprovider > gpurun_out/r2c45_pytest.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r2c45_pytest.log | cut -c1-300
timeout 300 python bench.py --gpus 2 --steps 100 --warmup 20 --no-extra > gpurun_out/r2c45_bench2.json 2> gpurun_out/r2c45_bench.err; python -c "
import json; d=json.load(open('gpurun_out/r2c45_bench2.json')); print('ddp 2 GPUs', d['ms_per_step'], d['e2e']['ms_per_step'], d['config']['sync'])" || tail -5 gpurun_out/r2c45_bench.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29671 bench/step_timeline.py --mode ddp > gpurun_out/r2c45_timeline.txt 2>&1; echo "rc=$?"; tail -75 gpurun_out/r2c45_timeline.txt
