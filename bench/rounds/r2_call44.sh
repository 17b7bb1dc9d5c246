#!/bin/bash
# round 2, call 44 (1 GPU): where the ~45 us between the kernel-group sum and the step go: overlap switches, same box
mkdir -p gpurun_out
for flags in "" "--no-overlap-wgrad" "--no-overlap-sgd" "--no-overlap-wgrad --no-overlap-sgd" "--no-pdl" ""; do
  timeout 600 python bench.py --steps 300 --warmup 30 --no-extra $flags > gpurun_out/r2c44_b.json 2> gpurun_out/r2c44_b.err
  python -c "
import json; d=json.load(open('gpurun_out/r2c44_b.json')); print('flags [$flags]:', d['ms_per_step'], d['e2e']['ms_per_step'], 'launches', d.get('gpu_launches'))" | tee -a gpurun_out/r2c44_ab.txt
done
