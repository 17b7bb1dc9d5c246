#!/bin/bash
mkdir -p gpurun_out
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/u2.json 2> gpurun_out/u2.err; echo rc=$?; cut -c1-250 gpurun_out/u2.json; tail -3 gpurun_out/u2.err
