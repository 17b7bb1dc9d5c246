#!/bin/bash
mkdir -p gpurun_out
run() { local name=$1; local n=$2; shift 2
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 200 --warmup 10 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/{n}.json") if l.startswith("{")][0]); print(n, d["ms_per_step"], d["value"], "e2e", d["e2e"]["ms_per_step"], d["config"].get("programmatic_dependent_launch"))
except Exception as e: print(n, "FAILED", e); print(open(f"gpurun_out/{n}.err").read()[-1500:])
PY
}
run y2_ddp_reference 2 --mode ddp
run y2_ddp_balanced 2 --mode ddp --bucket-layout balanced
run y2_ddp_perlayer 2 --mode ddp --bucket-layout per_layer
run y2_ddp_reference_bn2 2 --mode ddp --bn-fused 2
timeout 200 python bench.py --steps 200 --warmup 10 2>/dev/null | cut -c1-200
