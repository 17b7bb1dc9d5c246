#!/bin/bash
mkdir -p gpurun_out
run() { local name=$1; local n=$2; shift 2
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 200 --warmup 10 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/{n}.json") if l.startswith("{")][0]); print(n, d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d["config"]["sync"]["buckets_floats"])
except Exception as e: print(n, "FAILED", e); print(open(f"gpurun_out/{n}.err").read()[-1500:])
PY
}
run x2_reference 2 --mode ddp --bucket-layout reference
run x2_balanced 2 --mode ddp --bucket-layout balanced
run x2_balanced10 2 --mode ddp --bucket-layout balanced --bucket-mb 10
run x2_reference_b 2 --mode ddp --bucket-layout reference
