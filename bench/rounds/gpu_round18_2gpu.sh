#!/bin/bash
mkdir -p gpurun_out
run() { local name=$1; local n=$2; shift 2
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 100 --warmup 10 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  echo "$name: $(cut -c1-200 gpurun_out/$name.json)"; tail -2 gpurun_out/$name.err | grep -i error
}
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q --timeout 300 -k "ddp" > gpurun_out/tests18.log 2>&1; tail -3 gpurun_out/tests18.log
run t2_ddp_perlayer 2 --mode ddp
run t2_ddp_reference 2 --mode ddp --bucket-layout reference
run t2_ddp_perlayer_64 2 --mode ddp --comm-blocks 64
run t2_ddp_perlayer_16 2 --mode ddp --comm-blocks 16
run t2_allreduce 2 --mode allreduce
