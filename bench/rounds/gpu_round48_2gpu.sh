#!/bin/bash
mkdir -p gpurun_out
run() { local name=$1; local n=$2; shift 2
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 200 --warmup 10 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/{n}.json") if l.startswith("{")][0]); print(n, d["ms_per_step"], d["value"], "e2e", d["e2e"]["ms_per_step"], d["config"]["sync"]["comm_blocks"])
except Exception as e: print(n, "FAILED", e); print(open(f"gpurun_out/{n}.err").read()[-1500:])
PY
}
run v2_ddp_cb8 2 --mode ddp --comm-blocks 8
run v2_ddp_cb16 2 --mode ddp --comm-blocks 16
run v2_ddp_cb64 2 --mode ddp --comm-blocks 64
run v2_ddp_bal_cb16 2 --mode ddp --bucket-layout balanced --comm-blocks 16
run v2_ddp_pl_cb8 2 --mode ddp --bucket-layout per_layer --comm-blocks 8
