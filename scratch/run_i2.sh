#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dnn_gpu.py tests/test_encoder_gpu.py -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/i2_dnn_tests.log
timeout 600 python tools/resnet_bench.py --mode layers --relu-ch 2 --iters 3 > gpurun_out/i2_layers.json 2> gpurun_out/i2_layers.err; tail -c 900 gpurun_out/i2_layers.json; tail -2 gpurun_out/i2_layers.err
timeout 300 python tools/resnet_bench.py --mode conv --iters 3 > gpurun_out/i2_conv.json 2> gpurun_out/i2_conv.err; tail -c 600 gpurun_out/i2_conv.json
exit 0
