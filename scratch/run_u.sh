#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --resnet > gpurun_out/bench_2gpu_resnet.json 2> gpurun_out/bench_2gpu_resnet.err; echo "rc=$?"; tail -c 2500 gpurun_out/bench_2gpu_resnet.json; tail -3 gpurun_out/bench_2gpu_resnet.err
