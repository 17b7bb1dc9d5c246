#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest52.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest52.log
timeout 120 python bench.py --steps 200 --warmup 20 > gpurun_out/bench52.json 2> gpurun_out/bench52.err; echo "bench rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/bench52.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'], d['gpu_launches_per_step'])"
