#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "wgrad" 2>&1 | tail -2
for c in 4 1; do
echo "max cluster $c"
DDPB_WGRAD_CLUSTER=$c timeout 300 python bench/layer_graph_times.py 2>&1 | sed 's/fwd(conv+bn).*| wgrad/| wgrad/' | grep -v batch
done
run() { python bench.py --steps 200 --warmup 20 $2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', d['ms_per_step'], d['e2e']['ms_per_step'])"; }
DDPB_WGRAD_TAPS=1 run "wgrad tap-fused+cluster4"
DDPB_WGRAD_TAPS=0 run "wgrad one tap"
