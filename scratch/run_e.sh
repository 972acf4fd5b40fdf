#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file gpurun_out/launches_warm.csv python tools/boot_bench.py --mode synthetic --iters 1 --profile > gpurun_out/ncu_list_warm.log 2>&1
python tools/launch_summary.py gpurun_out/launches_warm.csv 2>/dev/null | head -12
