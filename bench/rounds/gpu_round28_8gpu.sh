#!/bin/bash
mkdir -p gpurun_out
run() { local name=$1; local n=$2; shift 2
  timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n --steps 200 --warmup 10 "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads([l for l in open(f"gpurun_out/{n}.json") if l.startswith("{")][0]); print(n, d["ms_per_step"], d["value"], "e2e", d["e2e"]["ms_per_step"], d["clocks"]["reasons"])
except Exception as e: print(n, "FAILED", e); print(open(f"gpurun_out/{n}.err").read()[-1500:])
PY
}
DDPB_TEST_WORLD=8 timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q --timeout 300 -k "multicast and ddp or p2p and gather_scatter or p2p and allreduce" > gpurun_out/tests28.log 2>&1
tail -3 gpurun_out/tests28.log
run f8_ddp_8 8 --mode ddp
run f8_ddp_4 4 --mode ddp
run f8_ddp_2 2 --mode ddp
timeout 200 python bench.py --steps 200 --warmup 10 > gpurun_out/f8_single_1.json 2>/dev/null; cut -c1-170 gpurun_out/f8_single_1.json
run f8_allreduce_8 8 --mode allreduce
run f8_gather_scatter_4 4 --mode gather_scatter
run f8_ddp_8_strong 8 --mode ddp --scaling strong
run f8_ddp_8_10mb 8 --mode ddp --bucket-mb 10
