#!/bin/bash
mkdir -p gpurun_out
timeout 300 python bench/layer_graph_times.py --bn-fused 2 > gpurun_out/layers36_f2.log 2>&1; cat gpurun_out/layers36_f2.log
timeout 300 python bench/layer_graph_times.py --bn-fused 0 2>&1 | grep -E "bn_bwd|totals" | sed 's/| wgrad.*//' > gpurun_out/layers36_f0.log; cat gpurun_out/layers36_f0.log
