#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_serial_gpu.py tests/test_keygen_gpu.py -x -q -s -m gpu 2>&1 | tail -40 | tee gpurun_out/serial_tests.log
