#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for i in 1 2; do
timeout 600 python tools/resnet_bench.py --mode layers --relu-ch 2 --iters 3 > gpurun_out/b2_layers_$i.json 2> gpurun_out/b2_layers.err; tail -c 900 gpurun_out/b2_layers_$i.json
done
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -5 | tee gpurun_out/b2_gpu_tests.log
exit 0
