#!/bin/bash
mkdir -p gpurun_out
run() { local name=$1; local n=$2; shift 2
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 100 --warmup 10 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/{n}.json") if l.startswith("{")][0]); print(n, d["ms_per_step"], d["e2e"]["ms_per_step"])
except Exception as e: print(n, "FAILED", e)
PY
}
run u2_perlayer_32 2 --mode ddp
run u2_perlayer_32_nofuse 2 --mode ddp --no-fuse-sgd
run u2_perlayer_16_nofuse 2 --mode ddp --no-fuse-sgd --comm-blocks 16
run u2_perlayer_8_nofuse 2 --mode ddp --no-fuse-sgd --comm-blocks 8
run u2_perlayer_64_nofuse 2 --mode ddp --no-fuse-sgd --comm-blocks 64
run u2_reference_32 2 --mode ddp --bucket-layout reference
run u2_reference_32_nofuse 2 --mode ddp --bucket-layout reference --no-fuse-sgd
run u2_allreduce 2 --mode allreduce
run u2_allreduce_nofuse 2 --mode allreduce --no-fuse-sgd
timeout 200 python bench.py --steps 100 --warmup 10 > gpurun_out/u1.json 2>/dev/null; cut -c1-200 gpurun_out/u1.json
