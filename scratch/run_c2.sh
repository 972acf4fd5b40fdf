#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
G=$(nvidia-smi -L | wc -l)
nvidia-smi topo -m 2>&1 | head -6
for n in 2 4 8; do
  [ $n -le $G ] || continue
  for extra in "--transport peer --graph" "--transport peer --no-gather --graph"; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n tools/limb_shard_bench.py $extra 2> gpurun_out/peer_shard_$n.err | grep config | tee -a gpurun_out/peer_shard_scaling.jsonl; grep -v "^\*\|OMP\|^$" gpurun_out/peer_shard_$n.err | tail -5
  done
done
exit 0
