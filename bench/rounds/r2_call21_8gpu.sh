#!/bin/bash
# round 2, call 21 (8 GPUs): multi-GPU tests at W=8 (NVLS and P2P paths) and W=3, bench (weak + extra configs), all-reduce sweep
mkdir -p gpurun_out
DDPB_TEST_WORLD=8 timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q -p no:cacheprovider -k "multicast or three_ranks or p2p or per_parameter" > gpurun_out/r2c21_pytest.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r2c21_pytest.log | cut -c1-300
timeout 600 python bench.py --gpus 8 --steps 30 --warmup 5 > gpurun_out/r2c21_ours8.json 2> gpurun_out/r2c21_ours8.err; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c21_ours8.json')); print(d['ms_per_step'], d['e2e']['ms_per_step']); [print(k, v['ms_per_step'], v['e2e_ms_per_step']) for k,v in d['extra_configs'].items()]"
DDPB_SWEEP_OUT=gpurun_out/r2c21_sweep8.json timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29655 bench/allreduce_sweep.py > gpurun_out/r2c21_sweep8.log 2>&1; echo "rc=$?"; tail -9 gpurun_out/r2c21_sweep8.log
