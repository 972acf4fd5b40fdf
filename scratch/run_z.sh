#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
G=$(nvidia-smi -L | wc -l)
for n in 2 4 8; do
  [ $n -le $G ] || continue
  for extra in "" "--no-gather"; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n tools/limb_shard_bench.py $extra 2> gpurun_out/limb_shard_$n.err | grep config | tee -a gpurun_out/limb_shard_scaling.jsonl; tail -3 gpurun_out/limb_shard_$n.err | grep -v "^\*\|OMP"
  done
done
