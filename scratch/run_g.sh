#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
timeout 300 python tools/boot_bench.py --mode synthetic --iters 5 2>&1 | tail -1 | tee gpurun_out/boot_syn.json
timeout 300 python tools/op_bench.py 2>&1 | tail -1 | tee gpurun_out/op_bench.json
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv python tools/boot_bench.py --mode synthetic --iters 1 --profile > gpurun_out/ncu_list.log 2>&1
python tools/launch_summary.py gpurun_out/launches.csv 2>/dev/null | head -16
