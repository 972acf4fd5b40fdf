#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
for extra in "" "--no-gather"; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/limb_shard_bench.py $extra 2> gpurun_out/limb_shard_2.err | grep config; tail -3 gpurun_out/limb_shard_2.err | grep -v "^\*\|OMP"
done
