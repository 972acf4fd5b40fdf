#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_seal_api_gpu.py -x -q -m gpu 2>&1 | tail -40 | tee gpurun_out/seal_tests.log
