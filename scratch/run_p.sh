#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_dnn_gpu.py tests/test_encoder_gpu.py -x -q -m gpu 2>&1 | tail -15 | tee gpurun_out/dnn_tests2.log
timeout 900 python tools/resnet_bench.py --mode layers --relu-ch 2 --iters 2 > gpurun_out/resnet_layers.json 2> gpurun_out/resnet_layers.err; echo "layers rc=$?"; cat gpurun_out/resnet_layers.json; tail -5 gpurun_out/resnet_layers.err
timeout 1200 python tools/resnet_bench.py --mode ours > gpurun_out/resnet_full.json 2> gpurun_out/resnet_full.err; echo "full rc=$?"; cat gpurun_out/resnet_full.json; tail -5 gpurun_out/resnet_full.err
