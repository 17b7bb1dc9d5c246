#!/bin/bash
mkdir -p gpurun_out
run() { python bench.py --steps 200 --warmup 20 $2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', d['ms_per_step'], d['e2e']['ms_per_step'])"; }
run "bn-fused 2" "--bn-fused 2"
run "bn-fused 1" "--bn-fused 1"
run "bn-fused 0" "--bn-fused 0"
run "bn-fused 2 again" "--bn-fused 2"
