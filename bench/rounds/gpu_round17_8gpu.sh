#!/bin/bash
# One 8-GPU call: sync-kernel tests at W=8, scaling benches (ours + reference), all-reduce sweep.
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
run() { # name gpus args...
  local name=$1; local n=$2; shift 2
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 100 --warmup 10 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  echo "$name: $(cut -c1-230 gpurun_out/$name.json)"
}
DDPB_TEST_WORLD=8 timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q --timeout 300 -k "multicast and ddp or p2p and gather_scatter or p2p and allreduce" > gpurun_out/tests17.log 2>&1
tail -4 gpurun_out/tests17.log
run s8_ours_ddp_8 8 --mode ddp
run s8_ours_ddp_4 4 --mode ddp
run s8_ours_ddp_2 2 --mode ddp
run s8_ours_allreduce_8 8 --mode allreduce
run s8_ours_gather_scatter_4 4 --mode gather_scatter
run s8_ours_ddp_8_strong 8 --mode ddp --scaling strong
run s8_ours_ddp_8_nomc 8 --mode ddp --no-multicast
run s8_ref_8 8 --impl reference
run s8_ref_4 4 --impl reference
run s8_ref_8_strong 8 --impl reference --scaling strong
DDPB_SWEEP_OUT=gpurun_out/allreduce_sweep_8.json timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29911 bench/allreduce_sweep.py > gpurun_out/sweep8.log 2> gpurun_out/sweep8.err
tail -3 gpurun_out/sweep8.err
