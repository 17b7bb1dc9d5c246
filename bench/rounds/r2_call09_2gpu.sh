#!/bin/bash
# round 2, call 9 (2 GPUs): multi-GPU tests again, ddp step timeline with the comm-stream kernels, bucket layouts
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c9_pytest.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/r2c9_pytest.log | cut -c1-300
for lay in reference per_layer balanced; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench/step_timeline.py --bucket-layout $lay > gpurun_out/r2c9_timeline_$lay.log 2>&1; grep -v "conv+bn fwd\|^\*\|^W1\|OMP" gpurun_out/r2c9_timeline_$lay.log | tail -50
done
run() { timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 "$@" 2> gpurun_out/r2c9_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['e2e']['ms_per_step'], d['loss_last'])" || tail -5 gpurun_out/r2c9_bench.err; }
run
run --bucket-layout per_layer
run --bucket-layout balanced
run --bucket-mb 10
