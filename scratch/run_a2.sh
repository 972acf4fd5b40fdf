#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_dnn_gpu.py tests/test_encoder_gpu.py -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/a2_dnn_tests.log
timeout 600 python tools/resnet_bench.py --mode layers --relu-ch 2 > gpurun_out/a2_layers.json 2> gpurun_out/a2_layers.err; tail -c 1500 gpurun_out/a2_layers.json
timeout 900 python tools/resnet_bench.py --mode ours > gpurun_out/a2_resnet.json 2> gpurun_out/a2_resnet.err; tail -c 1500 gpurun_out/a2_resnet.json
exit 0
