#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_dnn_gpu.py -x -q -s -m gpu -k "resnet_small" 2>&1 | tail -8 | tee gpurun_out/dnn_tests4.log
for l in 2 4 8; do
timeout 900 python tools/resnet_bench.py --mode layers --relu-ch 16 --iters 1 --lanes $l > gpurun_out/resnet_layers_t$l.json 2> gpurun_out/resnet_layers_t$l.err; echo "layers lanes=$l rc=$?"; python -c "
import json; print(json.load(open('gpurun_out/resnet_layers_t$l.json'))['relu_composite'])"; tail -3 gpurun_out/resnet_layers_t$l.err
done
