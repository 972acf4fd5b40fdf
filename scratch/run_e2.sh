#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
mkdir -p gpurun_out
G=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "peer_memory" 2>&1 | tail -3
for n in 2 4 8; do
  [ $n -le $G ] || continue
  for extra in "--transport peer --graph" "--transport peer --no-gather --graph"; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n tools/limb_shard_bench.py $extra 2> gpurun_out/peer_shard_$n.err | grep config | tee -a gpurun_out/peer_shard_scaling.jsonl; grep -v "^\*\|OMP\|^$" gpurun_out/peer_shard_$n.err | tail -5
  done
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $G --steps 10 --warmup 3 > gpurun_out/bench_${G}gpu.json 2> gpurun_out/bench_${G}gpu.err; tail -c 1500 gpurun_out/bench_${G}gpu.json; tail -3 gpurun_out/bench_${G}gpu.err
exit 0
