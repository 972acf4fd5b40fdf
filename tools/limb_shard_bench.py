#!/usr/bin/env python
"""Limb-sharded key switch (SURVEY.md 8e mode 2) under torchrun: ONE key switch at the reference's bootstrap chain
(N=2^16, |Q|=30, |P|=10, dnum=3) spread over the ranks' GPUs by RNS limb, NCCL broadcasts of the row blocks at the two
base-conversion boundaries.  Every rank holds the same synthetic ciphertext and key (same seed), checks the gathered result
against its own single-GPU key switch bit for bit, and the latency (CUDA events, MAX over ranks) is printed next to the
single-GPU latency.  TEST / BENCH infrastructure."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def parse(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--logn", type=int, default=16)
    ap.add_argument("--q", type=int, default=30)
    ap.add_argument("--p", type=int, default=10)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-gather", action="store_true", help="leave the result sharded (chained sharded operations)")
    ap.add_argument("--multicast", action="store_true", help="peer transport: rows stored once to the NVSwitch multicast address")
    ap.add_argument("--fused", action="store_true", help="peer transport with the stores to the peers inside the inverse transforms")
    ap.add_argument("--graph", action="store_true", help="also time the key switch replayed as one CUDA graph")
    ap.add_argument("--transport", default="nccl", choices=["nccl", "peer"],
                    help="row exchange: NCCL all-gather, or our own store kernel over NVLink peer memory")
    return ap.parse_args(argv)


def run(a, quiet=False):
    import torch
    import torch.distributed as dist
    import phantom_fhe_boot_b200 as fhe
    from phantom_fhe_boot_b200 import sharding
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    own_group = False
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", dev))
        own_group = True
    N, Q, P = 1 << a.logn, a.q, a.p
    primes = fhe.coeff_modulus_create(N, [60] + [59] * (Q - 1) + [60] * P)
    ctx = fhe.Context(a.logn, primes, P, device=dev)
    g = torch.Generator(device="cuda").manual_seed(1234)          # the same data on every rank

    def rnd(*shape):
        return torch.randint(0, 1 << 58, shape, dtype=torch.int64, device="cuda", generator=g)
    beta = ctx.beta(Q)
    evk = [rnd(2, Q + P, N) for _ in range(beta)]
    ct, c2 = rnd(2, Q, N), rnd(Q, N)
    want = ct.clone()
    ctx.keyswitch(Q, want, c2, evk)
    if a.transport == "peer" and world > 1:
        ks = sharding.PeerLimbShardedKeySwitch(ctx, Q, world=world, rank=rank, dist=dist, fused=a.fused, multicast=a.multicast)
    else:
        ks = sharding.LimbShardedKeySwitch(ctx, Q, world=world, rank=rank, dist=dist if world > 1 else None)
    out = torch.zeros_like(ct)
    for _ in range(3):  # repeated: buffer reuse across key switches must not race
        out.zero_()
        ks.run(c2, evk, acc=ct, acc_polys=0, out=out)
    torch.cuda.synchronize()
    ok = bool((out == want).all())

    def timed(fn):
        fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return sharding.max_over_ranks(e0.elapsed_time(e1) / a.iters, dist if world > 1 else None, "cuda")
    tmp = ct.clone()
    t_single = timed(lambda: ctx.keyswitch(Q, tmp, c2, evk))
    peer = a.transport == "peer" and world > 1   # the peer transport returns its arena's result block: no extra copy
    t_shard = timed(lambda: ks.run(c2, evk, acc=ct, acc_polys=0, out=None if peer else out, gather_out=not a.no_gather))
    phases = None
    if peer:  # per-phase device time on this rank (rank 0 reports), averaged over the iterations
        names, evs = [], []

        def mark(name):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            names.append(name)
            evs.append(e)
        dist.barrier()
        torch.cuda.synchronize()
        for _ in range(a.iters):
            mark("start")
            ks.run(c2, evk, acc=ct, acc_polys=0, gather_out=not a.no_gather, mark=mark)
        torch.cuda.synchronize()
        phases = {}
        for i in range(1, len(evs)):
            if names[i] != "start":
                phases[names[i]] = phases.get(names[i], 0.0) + evs[i - 1].elapsed_time(evs[i]) / a.iters
        phases = {k: round(v, 4) for k, v in phases.items()}
    t_graph = t_single_graph = None
    if a.graph:  # the launch sequence as ONE CUDA graph per key switch (the host issues ~30 small launches otherwise)
        def capture(fn):
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                fn()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                fn()
            return gr
        if peer or world == 1:
            gr = capture(lambda: ks.run(c2, evk, acc=ct, acc_polys=0, out=None if peer else out, gather_out=not a.no_gather))
            t_graph = timed(gr.replay)
        gs = capture(lambda: ctx.keyswitch(Q, tmp, c2, evk))
        t_single_graph = timed(gs.replay)
    if peer:  # still right after the back-to-back timed runs (no race between consecutive key switches)
        torch.cuda.synchronize()
        dist.barrier()
        res = ks.run(c2, evk, acc=ct, acc_polys=0, gather_out=True)
        torch.cuda.synchronize()
        ok = ok and bool((res == want).all())
    oks = [None] * world
    if world > 1:
        dist.all_gather_object(oks, ok)
    else:
        oks = [ok]
    result = None
    if rank == 0:
        result = ({"config": "key switch N=2^%d |Q|=%d |P|=%d dnum=%d, limbs over %d GPU(s)" % (a.logn, Q, P, beta, world),
                          "n_gpus": world, "bit_exact_vs_single_gpu_all_ranks": all(oks), "single_gpu_ms": round(t_single, 4),
                          "limb_sharded_ms_max_over_ranks": round(t_shard, 4), "speedup": round(t_single / t_shard, 3),
                          "cuda_graph": None if t_single_graph is None else {
                              "single_gpu_ms": round(t_single_graph, 4),
                              "limb_sharded_ms_max_over_ranks": None if t_graph is None else round(t_graph, 4)},
                          "phases_ms_rank0": phases, "result_gathered": not a.no_gather, "transport": (a.transport + ("" if a.transport == "nccl" else ("-fused" if a.fused else "-unfused") + ("-multicast" if getattr(ks, "multicast", False) else ""))) if world > 1 else "none",
                          "exchange_bytes_per_keyswitch": 8 * N * (Q + 2 * P + (0 if a.no_gather else 2 * Q))})
        if not quiet:
            print(json.dumps(result))
    ctx.close()
    if own_group:
        dist.destroy_process_group()
    return result


def main():
    run(parse())


if __name__ == "__main__":
    main()
