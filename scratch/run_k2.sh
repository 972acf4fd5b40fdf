#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for l in 4 6 8 12; do
timeout 600 python tools/resnet_bench.py --mode layers --relu-ch 16 --iters 2 --lanes $l > gpurun_out/k2_layers_$l.json 2> gpurun_out/k2_layers.err; python -c "
import json; d=json.load(open('gpurun_out/k2_layers_$l.json')); print($l, d['relu_composite'])"; tail -1 gpurun_out/k2_layers.err
done
exit 0
