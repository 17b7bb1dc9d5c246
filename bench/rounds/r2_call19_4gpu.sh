#!/bin/bash
# round 2, call 19 (4 GPUs): multi-GPU tests incl. W=3 (batch 85), bench both arms with extra_configs, all-reduce sweep
mkdir -p gpurun_out
DDPB_TEST_WORLD=4 timeout 1200 python -m pytest tests/test_multigpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c19_pytest.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/r2c19_pytest.log | cut -c1-300
timeout 600 python bench.py --gpus 4 --steps 30 --warmup 5 > gpurun_out/r2c19_ours4.json 2> gpurun_out/r2c19_ours4.err; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c19_ours4.json')); print(d['ms_per_step'], d['e2e']['ms_per_step']); [print(k, v['ms_per_step'], v['e2e_ms_per_step']) for k,v in d['extra_configs'].items()]"
timeout 600 python bench.py --gpus 4 --steps 30 --warmup 5 --impl reference > gpurun_out/r2c19_ref4.json 2> gpurun_out/r2c19_ref4.err; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c19_ref4.json')); print(d['ms_per_step']); [print(k, v['ms_per_step']) for k,v in d['extra_configs'].items()]"
DDPB_SWEEP_OUT=gpurun_out/r2c19_sweep4.json timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29655 bench/allreduce_sweep.py > gpurun_out/r2c19_sweep4.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/r2c19_sweep4.log
for v in "--mode gather_scatter" "--mode gather_scatter --no-multicast"; do timeout 300 python bench.py --gpus 4 --steps 30 --warmup 5 --no-extra $v 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['e2e']['ms_per_step'])"; done
