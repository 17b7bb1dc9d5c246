#!/bin/bash
# round 2, call 27 (2 GPUs): own VMM + multicast bootstrap, first-iteration single bucket, tightened backward oracle
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q -p no:cacheprovider -k "vmm or (p2p and ddp)" > gpurun_out/r2c27_pytest.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/r2c27_pytest.log | cut -c1-400
timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -x -q -p no:cacheprovider -k "precision_matched" > gpurun_out/r2c27_pytest2.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/r2c27_pytest2.log | cut -c1-300
DDPB_SYMM_BACKEND=vmm timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 --no-extra 2> gpurun_out/r2c27_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('vmm', d['ms_per_step'], d['e2e']['ms_per_step'], d['config']['sync'])" || tail -5 gpurun_out/r2c27_bench.err
