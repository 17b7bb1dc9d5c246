#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest44.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest44.log
timeout 300 python bench.py --steps 200 --warmup 20 > gpurun_out/bench44.json 2> gpurun_out/bench44.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/bench44.json
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
