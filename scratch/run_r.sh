#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 300 python tools/boot_bench.py --mode synthetic --slots 1024 --iters 5 > gpurun_out/boot_sparse.json 2> gpurun_out/boot_sparse.err; cat gpurun_out/boot_sparse.json; tail -3 gpurun_out/boot_sparse.err
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_sparse.csv python tools/boot_bench.py --mode synthetic --slots 1024 --iters 1 --profile > gpurun_out/ncu_list_sparse.log 2>&1
python tools/launch_summary.py gpurun_out/launches_sparse.csv > gpurun_out/launches_sparse_summary.json; cat gpurun_out/launches_sparse_summary.json | head -60
