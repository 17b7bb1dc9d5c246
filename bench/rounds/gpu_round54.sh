#!/bin/bash
mkdir -p gpurun_out
DDPB_TEST_EXPERIMENTAL=1 timeout 22 python -m pytest tests/test_experimental_gpu.py -m gpu -x -q -p no:cacheprovider > gpurun_out/pytest54.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest54.log | cut -c1-200
