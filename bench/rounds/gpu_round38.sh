#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q -k "wgrad or no_alloc" 2>&1 | tail -2
for d in 1 0; do
  DDPB_WGRAD_DEEP=$d timeout 300 python bench.py --steps 200 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('deep=$d', d['ms_per_step'], d['e2e']['ms_per_step'])"
done
