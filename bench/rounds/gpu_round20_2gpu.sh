#!/bin/bash
mkdir -p gpurun_out
run() { local name=$1; local n=$2; shift 2
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 100 --warmup 10 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/{n}.json") if l.startswith("{")][0]); print(n, d["ms_per_step"], d["e2e"]["ms_per_step"], d["config"]["sync"]["buckets_floats"])
except Exception as e: print(n, "FAILED", e); print(open(f"gpurun_out/{n}.err").read()[-1500:])
PY
}
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q --timeout 300 > gpurun_out/tests20.log 2>&1; tail -3 gpurun_out/tests20.log
run v2_ref 2 --mode ddp --bucket-layout reference
run v2_10mb 2 --mode ddp --bucket-mb 10
run v2_1mb 2 --mode ddp --bucket-mb 1
run v2_perlayer 2 --mode ddp
run v2_4mb 2 --mode ddp --bucket-mb 4
run v2_ref_64 2 --mode ddp --bucket-layout reference --comm-blocks 64
run v2_ref_16 2 --mode ddp --bucket-layout reference --comm-blocks 16
