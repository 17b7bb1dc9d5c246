#!/bin/bash
# round 2, call 3: full GPU test suite after the ABI / ADVICE changes
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/r2c3_pytest.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/r2c3_pytest.log | cut -c1-250
