#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
G=$(nvidia-smi -L | wc -l)
for n in 4; do
  [ $n -le $G ] || continue
  for extra in "--transport peer --no-gather --graph" "--transport peer --no-gather --graph --unfused"; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n tools/limb_shard_bench.py $extra 2> gpurun_out/peer_shard_$n.err | grep config | tee -a gpurun_out/peer_shard_scaling.jsonl; grep -v "^\*\|OMP\|^$" gpurun_out/peer_shard_$n.err | tail -5
  done
done
exit 0
