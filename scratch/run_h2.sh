#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python tools/resnet_bench.py --mode conv --iters 2 > gpurun_out/h2_conv.json 2> gpurun_out/h2_conv.err; tail -c 600 gpurun_out/h2_conv.json; tail -3 gpurun_out/h2_conv.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/h2_conv_launches.csv python tools/resnet_bench.py --mode conv --iters 1 > gpurun_out/h2_ncu.log 2>&1
python tools/launch_summary.py gpurun_out/h2_conv_launches.csv > gpurun_out/h2_conv_launches_summary.txt 2>&1; head -30 gpurun_out/h2_conv_launches_summary.txt
exit 0
