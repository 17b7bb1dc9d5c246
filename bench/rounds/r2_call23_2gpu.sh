#!/bin/bash
# round 2, call 23 (2 GPUs): ddp bucket layouts / comm CTAs again, with the release-store barrier and early per-bucket SGD
mkdir -p gpurun_out
run() { timeout 300 python bench.py --gpus 2 --steps 40 --warmup 10 --no-extra "$@" 2> gpurun_out/r2c23_bench.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$*', d['ms_per_step'], d['e2e']['ms_per_step'])" || tail -5 gpurun_out/r2c23_bench.err; }
run --bucket-layout balanced
run --bucket-layout balanced --comm-blocks 24
run --bucket-layout balanced --comm-blocks 32
run --bucket-layout balanced --comm-blocks 48
run --bucket-layout reference --comm-blocks 32
run --bucket-mb 5 --comm-blocks 32
run --bucket-mb 8 --comm-blocks 32
run --bucket-layout balanced --comm-blocks 32 --comm-threads 256
run --bucket-layout balanced --comm-blocks 32
