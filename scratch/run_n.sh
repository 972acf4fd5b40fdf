#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_dnn_gpu.py -x -q -s -m gpu 2>&1 | tail -60 | tee gpurun_out/dnn_tests.log
