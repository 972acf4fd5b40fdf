#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bsgs_kernel -s 4 -c 1 -o gpurun_out/m2_bsgs_compact -f python tools/resnet_bench.py --mode conv --iters 1 > gpurun_out/m2_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ntt_small_kernel -s 1 -c 1 -o gpurun_out/m2_ntt_sub -f python tools/resnet_bench.py --mode conv --iters 1 > gpurun_out/m2_ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decompose_lift_rescale -s 1 -c 1 -o gpurun_out/m2_decompose_lift -f python tools/resnet_bench.py --mode conv --iters 1 > gpurun_out/m2_ncu3.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/m2_ncu1.log
exit 0
