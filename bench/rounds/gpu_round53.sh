#!/bin/bash
mkdir -p gpurun_out
timeout 100 python bench.py --impl reference --steps 60 --warmup 5 > gpurun_out/ref53.json 2> gpurun_out/ref53.err; echo "ref rc=$?"; cut -c1-700 gpurun_out/ref53.json; tail -2 gpurun_out/ref53.err | cut -c1-200
