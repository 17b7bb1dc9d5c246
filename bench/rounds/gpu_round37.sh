#!/bin/bash
mkdir -p gpurun_out
for v in "" "--no-overlap-wgrad"; do
  timeout 300 python bench.py --steps 200 --warmup 20 $v 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['ms_per_step'], d['e2e']['ms_per_step'])"
done
DDPB_WGRAD_DIRECT=1 timeout 300 python bench.py --steps 200 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('wgrad_direct', d['ms_per_step'])"
DDPB_WGRAD_DIRECT=1 timeout 300 python bench.py --steps 200 --warmup 20 --no-overlap-wgrad 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('wgrad_direct no-overlap', d['ms_per_step'])"
