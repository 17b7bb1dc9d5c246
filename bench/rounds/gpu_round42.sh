#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q -k "wgrad or no_alloc or static" > gpurun_out/pytest42.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest42.log
timeout 300 python bench/layer_graph_times.py 2>&1 | sed 's/fwd(conv+bn).*| wgrad/| wgrad/' > gpurun_out/layers42.log; cat gpurun_out/layers42.log
run() { python bench.py --steps 200 --warmup 20 $2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', d['ms_per_step'], d['e2e']['ms_per_step'])"; }
DDPB_WGRAD_TAPS=1 run "wgrad tap-fused+cluster"
DDPB_WGRAD_TAPS=0 run "wgrad one tap"
