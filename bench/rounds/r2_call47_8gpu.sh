#!/bin/bash
# round 2, call 47 (8 GPUs): multi-GPU tests at W=8 (NVLS and P2P paths, W=3) and the bench (weak + extra configs) on the final kernels
mkdir -p gpurun_out
DDPB_TEST_WORLD=8 timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q -p no:cacheprovider -k "multicast or three_ranks or p2p or per_parameter" > gpurun_out/r2c47_pytest.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r2c47_pytest.log | cut -c1-300
timeout 600 python bench.py --gpus 8 --steps 60 --warmup 10 > gpurun_out/r2c47_ours8.json 2> gpurun_out/r2c47_ours8.err; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c47_ours8.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['value'], d['clocks']); [print(k, v['ms_per_step'], v['e2e_ms_per_step']) for k,v in d['extra_configs'].items()]"
