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
timeout 400 python -m pytest tests/test_multigpu.py -m gpu -q --timeout 300 -k "ddp" > gpurun_out/tests30.log 2>&1; tail -2 gpurun_out/tests30.log
run y2_reference_256 2 --mode ddp
run y2_reference_512 2 --mode ddp --comm-threads 512
run y2_balanced_256 2 --mode ddp --bucket-layout balanced
run y2_perlayer_256 2 --mode ddp --bucket-layout per_layer
run y2_reference_256_64 2 --mode ddp --comm-blocks 64
