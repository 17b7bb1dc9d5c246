#!/bin/bash
mkdir -p gpurun_out
echo "== alone" > gpurun_out/tests14.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 120 >> gpurun_out/tests14.log 2>&1
echo "== after the others" >> gpurun_out/tests14.log
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 --deselect tests/test_engine_gpu.py::test_matches_precision_matched_oracle >> gpurun_out/tests14.log 2>&1
cat gpurun_out/tests14.log | tail -20
