#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 1200 python tools/resnet_bench.py --mode ours --lanes 4 > gpurun_out/resnet_full_l4.json 2> gpurun_out/resnet_full_l4.err; echo "rc=$?"; cat gpurun_out/resnet_full_l4.json; tail -3 gpurun_out/resnet_full_l4.err
