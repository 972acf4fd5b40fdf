#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python tools/resnet_bench.py --mode ours --logn 12 --width 8 --in-ch 2 --base-ch 2 --blocks 1 --classes 3 > gpurun_out/resnet_small.json 2> gpurun_out/resnet_small.err; echo "small rc=$?"; cat gpurun_out/resnet_small.json; tail -5 gpurun_out/resnet_small.err
timeout 900 python tools/resnet_bench.py --mode layers --relu-ch 2 --iters 2 > gpurun_out/resnet_layers.json 2> gpurun_out/resnet_layers.err; echo "layers rc=$?"; cat gpurun_out/resnet_layers.json; tail -5 gpurun_out/resnet_layers.err
timeout 1200 python tools/resnet_bench.py --mode ours > gpurun_out/resnet_full.json 2> gpurun_out/resnet_full.err; echo "full rc=$?"; cat gpurun_out/resnet_full.json; tail -5 gpurun_out/resnet_full.err
