#!/bin/bash
mkdir -p gpurun_out
run() { python bench.py --steps 200 --warmup 20 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', d['ms_per_step'], d['e2e']['ms_per_step'])"; }
DDPB_FWD_FIN_IN_KERNEL=1 run "fwd-in-kernel=1 dgrad=0"
DDPB_FWD_FIN_IN_KERNEL=0 run "fwd-in-kernel=0 dgrad=0"
DDPB_FWD_FIN_IN_KERNEL=1 DDPB_DGRAD_FIN_IN_KERNEL=1 run "fwd-in-kernel=1 dgrad=1"
DDPB_FWD_FIN_IN_KERNEL=1 run "fwd-in-kernel=1 dgrad=0 (again)"
timeout 200 python bench/step_timeline.py > gpurun_out/timeline40.log 2>&1; cat gpurun_out/timeline40.log
