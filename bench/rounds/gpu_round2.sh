#!/bin/bash
# Engine integration tests, smoke, first bench lines (ours + reference) on one GPU.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -q --timeout 300 > gpurun_out/tests2.log 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
timeout 600 python bench.py --steps 50 --warmup 5 > gpurun_out/bench_ours_1.json 2> gpurun_out/bench_ours_1.err
timeout 600 python bench.py --impl reference --steps 50 --warmup 5 > gpurun_out/bench_ref_1.json 2> gpurun_out/bench_ref_1.err
tail -30 gpurun_out/tests2.log; tail -3 gpurun_out/smoke.log; cat gpurun_out/bench_ours_1.json; tail -5 gpurun_out/bench_ours_1.err; cat gpurun_out/bench_ref_1.json; tail -5 gpurun_out/bench_ref_1.err
