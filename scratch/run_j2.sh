#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python tools/resnet_bench.py --mode ours > gpurun_out/j2_resnet.json 2> gpurun_out/j2_resnet.err; tail -c 1200 gpurun_out/j2_resnet.json; tail -2 gpurun_out/j2_resnet.err
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4 | tee gpurun_out/j2_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
exit 0
