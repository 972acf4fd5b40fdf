#!/usr/bin/env python
"""ONE CKKS bootstrap on several GPUs: the linear-transform stages limb-sharded (b200fhe_boot_set_peer_shard,
sharding.PeerShardedBootstrap).  Launch with torchrun, one rank per GPU:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port 29641 \
      tools/boot_shard_bench.py [--logn 16 --levels-after 11 --special-primes 10 --iters 5]

Every rank builds the same synthetic context (same seed: the same keys, diagonals and input on every GPU), runs the
single-GPU bootstrap, then the sharded one, checks that every limb of the sharded result equals the single-GPU result on
EVERY rank, and times both with CUDA events (max over ranks).  Rank 0 prints one JSON line.
--virtual R: one process, one GPU, R virtual ranks (the row ranges without the exchange).  TEST / BENCH infrastructure."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def parse(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--logn", type=int, default=16)
    ap.add_argument("--levels-after", type=int, default=11)
    ap.add_argument("--special-primes", type=int, default=10)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--slots", type=int, default=0)
    ap.add_argument("--virtual", type=int, default=0)
    ap.add_argument("--no-graph", action="store_true")
    return ap.parse_args(argv)


def run(a, quiet=False):
    import torch
    import torch.distributed as dist
    import boot_bench
    import phantom_fhe_boot_b200 as fhe
    from phantom_fhe_boot_b200 import sharding
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx, boot, ct, info = boot_bench.synthetic_setup(a.logn, a.levels_after, a.special_primes, device=local, slots=a.slots, seed=1234)
    out = {"config": "N=2^%d |Q|=%d |P|=%d budget {2,2} slots=%d; uniform random keys/diagonals, same on every rank"
                     % (a.logn, info["Q"], a.special_primes, a.slots or info["N"] // 2), "world": world}

    def timed(fn, iters):
        fn()
        if world > 1:
            dist.barrier()
        ms = boot_bench.ev_time(torch, fn, iters)
        return sharding.max_over_ranks(ms, dist if world > 1 else None, "cuda")

    want = boot.eval(ct).tensor().clone()
    torch.cuda.synchronize()
    out["single_gpu_ms"] = round(timed(lambda: boot.eval(ct), a.iters), 3)
    if not a.no_graph:   # the same bootstrap replayed as a CUDA graph (b200fhe_boot_graph_capture)
        g = boot.capture(ct)
        got = g.launch().tensor()
        torch.cuda.synchronize()
        out["single_gpu_graph_bit_exact"] = bool((got == want).all())
        out["single_gpu_graph_ms"] = round(timed(lambda: g.launch(), a.iters), 3)
        g.close()
    if a.virtual > 1:
        boot.set_peer_shard(a.virtual, virtual_ranks=True)
        got = boot.eval(ct).tensor()
        torch.cuda.synchronize()
        out["virtual_ranks"] = a.virtual
        out["bit_exact_vs_single_gpu_all_ranks"] = bool((got == want).all())
        out["virtual_ms"] = round(timed(lambda: boot.eval(ct), a.iters), 3)
        boot.set_peer_shard(1)
    elif world > 1:
        sh = sharding.PeerShardedBootstrap(boot, world, rank, dist)
        n0 = fhe.launch_count()
        got = boot.eval(ct).tensor()
        torch.cuda.synchronize()
        out["launches_per_bootstrap_per_rank"] = fhe.launch_count() - n0
        ok = torch.tensor([1 if bool((got == want).all()) else 0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        out["bit_exact_vs_single_gpu_all_ranks"] = bool(ok.item())
        out["limb_sharded_ms"] = round(timed(lambda: boot.eval(ct), a.iters), 3)
        # host time to ISSUE one sharded bootstrap (the call returns before the GPU finishes): launch-bound when close to the above
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        boot.eval(ct)
        out["host_issue_ms_rank0"] = round((time.perf_counter() - t0) * 1e3, 2)
        torch.cuda.synchronize()
        # a second result after the timed runs (both arena halves reused many times)
        got = boot.eval(ct).tensor()
        torch.cuda.synchronize()
        ok = torch.tensor([1 if bool((got == want).all()) else 0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        out["bit_exact_after_reuse"] = bool(ok.item())
        if not a.no_graph:
            g = boot.capture(ct)
            got = g.launch().tensor()
            torch.cuda.synchronize()
            ok = torch.tensor([1 if bool((got == want).all()) else 0], device="cuda")
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            out["limb_sharded_graph_bit_exact_all_ranks"] = bool(ok.item())
            out["limb_sharded_graph_ms"] = round(timed(lambda: g.launch(), a.iters), 3)
            got = g.launch().tensor()
            torch.cuda.synchronize()
            ok = torch.tensor([1 if bool((got == want).all()) else 0], device="cuda")
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            out["limb_sharded_graph_bit_exact_after_replays"] = bool(ok.item())
            g.close()
        out["speedup"] = round(out["single_gpu_ms"] / min(out["limb_sharded_ms"], out.get("limb_sharded_graph_ms", 1e9)), 3)
        out["arena_MiB_per_rank"] = round(boot.peer_arena_bytes() / 2 ** 20, 1)
        out["transport"] = "peer memory (b200fhe_peer_exchange store kernel, no NCCL on the data path)"
        sh.close()
    # give the 70 GB of keys, diagonals and pools back before the caller goes on
    del want, ct
    boot.close()
    ctx.close()
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    if rank == 0 and not quiet:
        print(json.dumps(out), flush=True)
    return out


if __name__ == "__main__":
    run(parse())
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()
