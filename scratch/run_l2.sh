#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for l in 8 4; do
timeout 900 python tools/resnet_bench.py --mode ours --lanes $l > gpurun_out/l2_resnet_$l.json 2> gpurun_out/l2_resnet.err; tail -c 700 gpurun_out/l2_resnet_$l.json; tail -1 gpurun_out/l2_resnet.err
done
exit 0
