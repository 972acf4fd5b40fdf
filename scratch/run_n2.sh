#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_bootstrap_gpu.py tests/test_dnn_gpu.py -x -q -m gpu -k "not reference_configuration" 2>&1 | tail -3
timeout 300 python tools/resnet_bench.py --mode conv --iters 3 2> gpurun_out/n2.err | tail -c 400
timeout 600 python tools/boot_bench.py --mode synthetic --iters 5 2>> gpurun_out/n2.err | tail -c 500
timeout 600 python tools/boot_bench.py --mode synthetic --iters 5 --slots 1024 2>> gpurun_out/n2.err | tail -c 500
tail -2 gpurun_out/n2.err
exit 0
