#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "limb_sharded" 2>&1 | tail -15
timeout 300 python tools/limb_shard_bench.py > gpurun_out/limb_shard_1.json 2> gpurun_out/limb_shard_1.err; cat gpurun_out/limb_shard_1.json; tail -3 gpurun_out/limb_shard_1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/limb_shard_bench.py > gpurun_out/limb_shard_2.json 2> gpurun_out/limb_shard_2.err; cat gpurun_out/limb_shard_2.json; tail -3 gpurun_out/limb_shard_2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 tools/limb_shard_bench.py --no-gather > gpurun_out/limb_shard_2ng.json 2> gpurun_out/limb_shard_2ng.err; cat gpurun_out/limb_shard_2ng.json; tail -3 gpurun_out/limb_shard_2ng.err
