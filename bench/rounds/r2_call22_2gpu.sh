#!/bin/bash
# round 2, call 22 (2 GPUs): release-store peer barrier (tests, small-message sweep, ddp bench) and the wgrad deep ring A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c22_pytest.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/r2c22_pytest.log | cut -c1-300
DDPB_SWEEP_OUT=gpurun_out/r2c22_sweep2.json timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 bench/allreduce_sweep.py > gpurun_out/r2c22_sweep2.log 2>&1; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/r2c22_sweep2.json'))
for r in d['rows']: print(r['what'], round(r['ours_best_us'],1), round(r['nccl_us'],1))
print(d['per_parameter'])"
run() { timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 --no-extra "$@" 2> gpurun_out/r2c22_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['e2e']['ms_per_step'])" || tail -5 gpurun_out/r2c22_bench.err; }
run; run
for d in 1 0 1 0; do DDPB_WGRAD_DEEP=$d timeout 200 python bench.py --steps 40 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('deep=$d', d['ms_per_step'], d['e2e']['ms_per_step'])"; done
for d in 1 0; do DDPB_WGRAD_DEEP=$d timeout 200 python bench/layer_graph_times.py 2>&1 | grep -E "^L[5-7]|totals" | sed 's/fwd(conv+bn) *[0-9.]* | bn_bwd *[0-9.]* |//'; done
