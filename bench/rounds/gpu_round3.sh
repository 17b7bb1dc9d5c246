#!/bin/bash
# 2-GPU: gradient diagnostics (1 GPU), sync-kernel tests, 2-GPU bench lines ours + reference.
mkdir -p gpurun_out
timeout 200 python bench/diag_grads.py 64 > gpurun_out/diag64.log 2>&1
timeout 200 python bench/diag_grads.py 24 > gpurun_out/diag24.log 2>&1
NCCL_DEBUG=WARN timeout 1500 python -m pytest tests/test_multigpu.py -m gpu -q --timeout 600 > gpurun_out/tests3.log 2>&1
for mode in ddp allreduce gather_scatter; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 50 --warmup 5 --mode $mode > gpurun_out/bench_ours_2_$mode.json 2> gpurun_out/bench_ours_2_$mode.err
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29556 bench.py --gpus 2 --steps 50 --warmup 5 --impl reference > gpurun_out/bench_ref_2.json 2> gpurun_out/bench_ref_2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29557 bench.py --gpus 2 --steps 50 --warmup 5 --mode ddp --no-multicast > gpurun_out/bench_ours_2_ddp_nomc.json 2> gpurun_out/bench_ours_2_ddp_nomc.err
head -3 gpurun_out/diag64.log; tail -15 gpurun_out/tests3.log; cat gpurun_out/bench_ours_2_*.json gpurun_out/bench_ref_2.json | cut -c1-400; tail -3 gpurun_out/*2*.err
