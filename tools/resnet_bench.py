#!/usr/bin/env python
"""Encrypted ResNet-20 inference on the reference's configuration (Resnet/resnet.cu: N=2^16, |Q|=30 (60 + 29x59 bit),
|P|=10, level budget {2,2}, 32x32x3 image on 1024 sparse slots, scale 2^59) with SYNTHETIC weights and image (the
trained weights / CIFAR file are not in the reference tree).

  --mode ours       our whole stack, own keys (ChaCha20 key generator), own diagonals: full network (or --blocks /
                    --base-ch for a smaller one); per-layer-class times, decrypted result against the plain numpy network.
  --mode layers     ONE convolution (base_ch -> base_ch on the full image) and ONE ReluComposite on --relu-ch channels,
                    ours next to the reference's GPU build (oracle/_ref) on the reference's keys: bit-exact check + times.
Under torchrun each rank runs its own image on its own GPU (batch sharding, no data-path collective); rank 0 prints
the max-over-ranks time.  Prints one JSON object.  TEST / BENCH infrastructure."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import resnet_model  # noqa: E402

SCALE = 2.0 ** 59


def chain_bits(levels_after, nP):
    return [60] + [59] * (levels_after + 18) + [60] * nP


def run_ours(a):
    import torch
    import phantom_fhe_boot_b200 as fhe
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(dev)
    own_group = False
    if world > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device("cuda", dev))
            own_group = True
    N, W = 1 << a.logn, a.width
    bits = chain_bits(a.levels_after, a.np)
    out = {"mode": "ours", "config": "N=2^%d |Q|=%d |P|=%d image %dx%dx%d base_ch=%d blocks=%d classes=%d slots=%d" % (
        a.logn, len(bits) - a.np, a.np, W, W, a.in_ch, a.base_ch, a.blocks, a.classes, W * W), "n_gpus": world}
    t0 = time.time()
    primes = fhe.coeff_modulus_create(N, bits)
    ctx = fhe.Context(a.logn, primes, a.np, device=dev)
    boot = fhe.Bootstrapper(ctx, (2, 2), num_slots=W * W)
    enc = fhe.Encoder(ctx)
    boot.precompute(enc)
    # --one-image: ONE image on all ranks (channels of the ReLU / bootstrap layers dealt to the ranks): same keys, same image
    krank = 0 if a.one_image else rank
    kg = fhe.KeyGenerator(ctx, seed=bytes((7 * i + krank) % 256 for i in range(32)))
    ev = boot.ev
    rlk = kg.relin_key()
    ev.set_relin_key(rlk.digits)
    brots = boot.rotation_indices()
    lrots = fhe.resnet20_rotation_indices(W)
    keys = kg.rotation_keys(sorted(set(brots) | set(lrots)), conj=True)
    for k, key in keys.items():
        ev.set_galois_key(k, key.digits)
    pk = kg.public_key()
    dnn = fhe.DNN(ev, enc, SCALE)
    dnn.set_lanes(boot, a.lanes)
    dnn.set_batch(a.batch)
    out["lanes"] = a.lanes
    out["batch"] = a.batch
    torch.cuda.synchronize()
    out["setup_s"] = round(time.time() - t0, 2)
    out["galois_keys"] = len(keys)
    free, total = torch.cuda.mem_get_info()
    out["hbm_used_GiB_after_setup"] = round((total - free) / 2 ** 30, 1)

    cfg = dict(in_ch=a.in_ch, base_ch=a.base_ch, blocks=a.blocks, classes=a.classes)
    net, flat = resnet_model.synthetic_params(a.seed, **cfg)
    img = np.random.default_rng(a.seed + 100 + krank).uniform(-1, 1, (W, W, a.in_ch))
    x = dnn.enc_tensor(kg, pk, img)
    shard = None
    if a.one_image and world > 1:
        from phantom_fhe_boot_b200 import sharding
        widest = 4 * a.base_ch
        shard = sharding.PeerShardedDNN(dnn, world, rank, dist, 2 * widest * 2 * 16 * N * 8)
        out["mode"] = "ours, one image on %d GPUs (channels of ReluComposite / BootStrap and the input channels of Conv dealt to the ranks)" % world
    n0 = fhe.launch_count()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.time()
    y, ms = fhe.resnet20_infer(dnn, boot, x, flat, base_ch=a.base_ch, blocks=a.blocks, classes=a.classes, timings=True)
    torch.cuda.synchronize()
    dt = time.time() - t0
    if world > 1:
        t = torch.tensor([dt], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    out["inference_s"] = round(dt, 3)
    out["images_per_s"] = round((1 if a.one_image else world) / dt, 4)
    if a.one_image:
        # every rank must hold the same result limbs: compare a checksum of the output ciphertexts across the ranks
        chk = torch.zeros(1, dtype=torch.int64, device="cuda")
        for c in y.cts:
            chk += c.tensor().sum()
        out["output_checksum"] = int(chk.item())
        if world > 1:
            lo, hi = chk.clone(), chk.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            out["same_limbs_on_every_rank"] = bool(lo.item() == hi.item())
    out["launches"] = int(fhe.launch_count() - n0)
    out["layer_ms"] = {k: round(v, 1) for k, v in ms.items()}
    relu_ch = a.base_ch * (1 + 2 * a.blocks) + 2 * a.base_ch * 2 * a.blocks + 4 * a.base_ch * 2 * a.blocks
    out["bootstraps"] = 3 * relu_ch
    out["pool_GiB"] = round(ev.pool_bytes() / 2 ** 30, 2)
    # values: entry u >= 1 of the result is (weight[0][u] + ...) as the reference computes it (see AvgPoolFullCon in
    # include/b200fhe.h); check the network itself through the pooled features: sum_k fc_w[u][k] * pooled[k] + bias
    feat, pooled = resnet_model.forward_plain(net, img)
    res = dnn.dec_tensor(kg, fhe.TensorCT(y.cts, 1, 0, len(y.cts)))[0, 0, :]
    want = np.array([net["fc_w"][0, u] * pooled[u] for u in range(len(y.cts))])
    for u in range(1, a.classes):
        want[u] += net["fc_w"][u, :] @ pooled
    want[:a.classes] += net["fc_b"]
    out["max_abs_err_vs_plain"] = float(np.abs(res - want).max())
    out["levels_out"] = int(y.cts[0].desc().size_Ql)
    if shard is not None:
        shard.close()
    if rank == 0:
        print(json.dumps(out))
    if own_group:
        dist.destroy_process_group()
    return out


def run_conv(a):
    """our convolution alone on our own keys (no reference, no bootstrap): what a profiler run wants"""
    import torch
    import phantom_fhe_boot_b200 as fhe
    N, W = 1 << a.logn, a.width
    bits = chain_bits(a.levels_after, a.np)
    primes = fhe.coeff_modulus_create(N, bits)
    ctx = fhe.Context(a.logn, primes, a.np)
    boot = fhe.Bootstrapper(ctx, (2, 2), num_slots=W * W)     # only for its evaluator; no diagonals are computed
    enc = fhe.Encoder(ctx)
    kg = fhe.KeyGenerator(ctx, seed=bytes(range(32)))
    ev = boot.ev
    ev.set_relin_key(kg.relin_key().digits)
    keys = kg.rotation_keys(sorted(set(fhe.DNN.add_rotation_indices([], W, 3, 0))), conj=False)
    for k, key in keys.items():
        ev.set_galois_key(k, key.digits)
    pk = kg.public_key()
    dnn = fhe.DNN(ev, enc, SCALE)
    rng = np.random.default_rng(a.seed)
    img = rng.uniform(-1, 1, (W, W, a.base_ch))
    wt = rng.uniform(-0.15, 0.15, (3, 3, a.base_ch, a.base_ch))
    x = dnn.enc_tensor(kg, pk, img)
    out = {"mode": "conv", "config": "N=2^%d |Q|=%d |P|=%d conv %d->%d 3x3 on %dx%d, slots=%d" % (
        a.logn, len(bits) - a.np, a.np, a.base_ch, a.base_ch, W, W, W * W)}
    for name in ("conv_fresh_input", "conv_noise2_input"):
        y = dnn.conv(x, wt, 1)
        torch.cuda.synchronize()
        n0, t0 = fhe.launch_count(), time.time()
        for _ in range(a.iters):
            y = dnn.conv(x, wt, 1)
        torch.cuda.synchronize()
        out[name] = {"ours_ms": round((time.time() - t0) / a.iters * 1e3, 1), "launches": int((fhe.launch_count() - n0) / a.iters)}
        x = y
    print(json.dumps(out))
    return out


def run_full_vs_reference(a):
    """The WHOLE network on the reference's GPU build and on ours, same keys, same encrypted image, same synthetic weights:
    final ciphertext limbs compared, both decrypted results compared with the plain float network, both timed.
    (The reference's own driver, Resnet/models/model_resnet20.cu, needs cnpy / zlib / weight files that are not in its
    tree; this walks the same layer sequence through its DNN class.)"""
    import torch
    import phantom_fhe_boot_b200 as fhe
    import refgpu
    R = refgpu.lib()
    N, W = 1 << a.logn, a.width
    bits = chain_bits(a.levels_after, a.np)
    out = {"mode": "full_vs_reference", "config": "ResNet-20 shape: in %d, base %d, %d blocks per stage, %d classes, %dx%d image; "
           "N=2^%d |Q|=%d |P|=%d slots=%d" % (a.in_ch, a.base_ch, a.blocks, a.classes, W, W, a.logn, len(bits) - a.np, a.np, W * W)}
    ref = refgpu.Ref(a.logn, bits, a.np)
    rb = refgpu.RefBoot(ref, (2, 2), SCALE, num_slots=W * W, keygen=True)
    ctx = fhe.Context(a.logn, ref.primes, a.np)
    boot = fhe.Bootstrapper(ctx, (2, 2), num_slots=W * W)
    enc = fhe.Encoder(ctx)
    boot.precompute(enc)
    rb.install_keys(boot.ev, boot.rotation_indices())
    rots = fhe.resnet20_rotation_indices(W)
    rd = refgpu.RefDnn(rb, SCALE, rots)
    dnn = fhe.DNN(boot.ev, enc, SCALE)
    rd.install_keys(dnn)
    dnn.set_lanes(boot, a.lanes)
    dnn.set_batch(a.batch)
    cfg = dict(in_ch=a.in_ch, base_ch=a.base_ch, blocks=a.blocks, classes=a.classes)
    net, flat = resnet_model.synthetic_params(a.seed, **cfg)
    img = np.random.default_rng(a.seed + 100).uniform(-1, 1, (W, W, a.in_ch))
    rt = rd.enc_tensor(img)
    cts = []
    for i in range(R.rref_tensor_size(rt)):
        rct = R.rref_tensor_ct(rt, i)
        cts.append(boot.ev.import_ct(ref.ct_tensor(rct), R.rref_ct_limbs(rct), polys=R.rref_ct_size(rct),
                                     scale=R.rref_ct_scale(rct), noise=R.rref_ct_noise(rct)))
    ot = fhe.TensorCT(cts, R.rref_tensor_width(rt), R.rref_tensor_slotstr(rt), R.rref_tensor_num_ch(rt))

    # ours
    torch.cuda.synchronize()
    n0 = fhe.launch_count()
    t0 = time.time()
    y, ms = fhe.resnet20_infer(dnn, boot, ot, flat, base_ch=a.base_ch, blocks=a.blocks, classes=a.classes, timings=True)
    torch.cuda.synchronize()
    out["ours_s"] = round(time.time() - t0, 2)
    out["ours_launches"] = int(fhe.launch_count() - n0)
    out["ours_layer_ms"] = {k: round(v, 1) for k, v in ms.items()}

    # the reference, layer by layer in the same order (csrc/resnet20.cu = Resnet/models/model_resnet20.cu)
    layer_s = {"conv": 0.0, "bn_add": 0.0, "relu_composite": 0.0, "avgpool_fullcon": 0.0}

    def timed(key, f, *args):
        ref.sync()
        t = time.time()
        r = f(*args)
        ref.sync()
        layer_s[key] += time.time() - t
        return r

    def r_relu(t):
        return timed("relu_composite", lambda: rd.t(R.rref_dnn_relu_composite, t))
    t0 = time.time()
    x = r_relu(timed("bn_add", rd.batch_norm, timed("conv", rd.conv, rt, net["conv1"], 1), *net["bn1"]))
    for blk in net["blocks"]:
        skip = x
        yb = r_relu(timed("bn_add", rd.batch_norm, timed("conv", rd.conv, x, blk["w1"], 2 if blk["down"] else 1), *blk["bn1"]))
        yb = timed("bn_add", rd.batch_norm, timed("conv", rd.conv, yb, blk["w2"], 1), *blk["bn2"])
        if blk["down"]:
            skip = timed("bn_add", rd.batch_norm, timed("conv", rd.conv, skip, blk["ws"], 2), *blk["bns"])
        yb = timed("bn_add", lambda: rd.t(R.rref_dnn_add, yb, skip))
        x = r_relu(yb)
    rres = timed("avgpool_fullcon", rd.avgpool_fullcon, x, net["fc_w"], net["fc_b"])
    ref.sync()
    out["reference_s"] = round(time.time() - t0, 2)
    out["reference_layer_s"] = {k: round(v, 2) for k, v in layer_s.items()}
    out["speedup"] = round(out["reference_s"] / out["ours_s"], 2)

    # limbs, then values against the plain float network
    same = len(y.cts) == R.rref_tensor_size(rres)
    for i, ct in enumerate(y.cts):
        same = same and bool((ct.tensor().reshape(-1) == ref.ct_tensor(R.rref_tensor_ct(rres, i)).reshape(-1)).all())
    out["final_limbs_bit_exact"] = bool(same)
    feat, pooled = resnet_model.forward_plain(net, img)
    want = np.array([net["fc_w"][0, u] * pooled[u] for u in range(len(y.cts))])
    for u in range(1, a.classes):
        want[u] += net["fc_w"][u, :] @ pooled
    want[:a.classes] += net["fc_b"]
    kg = fhe.KeyGenerator(ctx, seed=bytes(range(32)))
    sk = torch.empty((len(bits), N), dtype=torch.int64, device="cuda")
    refgpu.ck(R.rref_secret_key(ref.h, sk.data_ptr()))
    kg.set_secret(sk)
    got = dnn.dec_tensor(kg, fhe.TensorCT(y.cts, 1, 0, len(y.cts)))[0, 0, :]
    out["ours_max_abs_err_vs_plain"] = float(np.abs(got - want).max())
    out["argmax_plain"], out["argmax_ours"] = int(np.argmax(want[:a.classes])), int(np.argmax(got[:a.classes]))
    out["logit_spread_plain"] = float(want[:a.classes].max() - np.sort(want[:a.classes])[-2])
    print(json.dumps(out))
    return out


def run_layers(a):
    import torch
    import phantom_fhe_boot_b200 as fhe
    import refgpu
    R = refgpu.lib()
    N, W = 1 << a.logn, a.width
    bits = chain_bits(a.levels_after, a.np)
    out = {"mode": "layers", "config": "N=2^%d |Q|=%d |P|=%d conv %d->%d 3x3 on %dx%d, ReluComposite on %d channels, slots=%d" % (
        a.logn, len(bits) - a.np, a.np, a.base_ch, a.base_ch, W, W, a.relu_ch, W * W)}
    ref = refgpu.Ref(a.logn, bits, a.np)
    rb = refgpu.RefBoot(ref, (2, 2), SCALE, num_slots=W * W, keygen=True)
    ctx = fhe.Context(a.logn, ref.primes, a.np)
    boot = fhe.Bootstrapper(ctx, (2, 2), num_slots=W * W)
    enc = fhe.Encoder(ctx)
    boot.precompute(enc)
    brots = boot.rotation_indices()
    rb.install_keys(boot.ev, brots)
    rots = fhe.DNN.add_rotation_indices([], W, 3, 0)
    rd = refgpu.RefDnn(rb, SCALE, rots)
    dnn = fhe.DNN(boot.ev, enc, SCALE)
    rd.install_keys(dnn)
    dnn.set_lanes(boot, a.lanes)
    dnn.set_batch(a.batch)
    out["lanes"] = a.lanes
    out["batch"] = a.batch
    rng = np.random.default_rng(a.seed)
    img = rng.uniform(-1, 1, (W, W, a.base_ch))
    wt = rng.uniform(-0.15, 0.15, (3, 3, a.base_ch, a.base_ch))
    rt = rd.enc_tensor(img)

    def ours_of(rt_):
        cts = []
        for i in range(R.rref_tensor_size(rt_)):
            rct = R.rref_tensor_ct(rt_, i)
            cts.append(boot.ev.import_ct(ref.ct_tensor(rct), R.rref_ct_limbs(rct), polys=R.rref_ct_size(rct),
                                         scale=R.rref_ct_scale(rct), noise=R.rref_ct_noise(rct)))
        return fhe.TensorCT(cts, R.rref_tensor_width(rt_), R.rref_tensor_slotstr(rt_), R.rref_tensor_num_ch(rt_))

    def equal(o, r):
        ref.sync()
        torch.cuda.synchronize()
        ok = len(o.cts) == R.rref_tensor_size(r)
        for i, ct in enumerate(o.cts):
            ok = ok and bool((ct.tensor().reshape(-1) == ref.ct_tensor(R.rref_tensor_ct(r, i)).reshape(-1)).all())
        return ok

    def wall(fn, iters):
        fn()
        torch.cuda.synchronize()
        ref.sync()
        t0 = time.time()
        for _ in range(iters):
            r = fn()
        torch.cuda.synchronize()
        ref.sync()
        return (time.time() - t0) / iters * 1e3, r

    ot = ours_of(rt)
    # a NoiseScaleDeg-2 input like every convolution after the first sees: one BatchNorm-free conv output
    t_ref, r1 = wall(lambda: rd.conv(rt, wt, 1), 1)
    t_our, o1 = wall(lambda: dnn.conv(ot, wt, 1), a.iters)
    out["conv_fresh_input"] = {"reference_gpu_ms": round(t_ref, 1), "ours_ms": round(t_our, 1), "bit_exact": equal(o1, r1),
                               "speedup": round(t_ref / t_our, 2)}
    t_ref, r2 = wall(lambda: rd.conv(r1, wt, 1), 1)
    t_our, o2 = wall(lambda: dnn.conv(o1, wt, 1), a.iters)
    out["conv_noise2_input"] = {"reference_gpu_ms": round(t_ref, 1), "ours_ms": round(t_our, 1), "bit_exact": equal(o2, r2),
                                "speedup": round(t_ref / t_our, 2)}
    # ReluComposite on the first relu_ch channels
    arr = (refgpu.vp * a.relu_ch)(*[R.rref_tensor_ct(r1, i) for i in range(a.relu_ch)])
    r1s = R.rref_tensor_new(arr, a.relu_ch, W, 0, a.relu_ch)
    o1s = fhe.TensorCT(o1.cts[:a.relu_ch], W, 0, a.relu_ch)
    t_ref, r3 = wall(lambda: rd.t(R.rref_dnn_relu_composite, r1s), 1)
    t_our, o3 = wall(lambda: dnn.relu_composite(o1s, boot), a.iters)
    out["relu_composite"] = {"channels": a.relu_ch, "reference_gpu_ms": round(t_ref, 1), "ours_ms": round(t_our, 1),
                             "bit_exact": equal(o3, r3), "speedup": round(t_ref / t_our, 2),
                             "ours_ms_per_bootstrap_incl_sign": round(t_our / (3 * a.relu_ch), 2)}
    if a.sweep:   # the same ReluComposite under other (lanes, batch) settings, bit-compared with the reference's result
        out["relu_sweep"] = []
        for pair in a.sweep.split(","):
            l, b = (int(v) for v in pair.split(":"))
            dnn.set_lanes(boot, l)
            dnn.set_batch(b)
            t, o = wall(lambda: dnn.relu_composite(o1s, boot), a.iters)
            out["relu_sweep"].append({"lanes": l, "batch": b, "ms": round(t, 1), "bit_exact": equal(o, r3)})
    print(json.dumps(out))
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="ours", choices=["ours", "layers", "conv", "full_vs_reference"])
    ap.add_argument("--logn", type=int, default=16)
    ap.add_argument("--levels-after", type=int, default=11)
    ap.add_argument("--np", type=int, default=10)
    ap.add_argument("--width", type=int, default=32)
    ap.add_argument("--in-ch", type=int, default=3)
    ap.add_argument("--base-ch", type=int, default=16)
    ap.add_argument("--blocks", type=int, default=3)
    ap.add_argument("--classes", type=int, default=10)
    ap.add_argument("--relu-ch", type=int, default=2)
    ap.add_argument("--iters", type=int, default=2)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--lanes", type=int, default=4)
    ap.add_argument("--sweep", default="", help="layers mode: more (lanes:batch) settings for the ReluComposite, e.g. 4:1,2:4,0:8")
    ap.add_argument("--one-image", action="store_true",
                    help="under torchrun: ONE image on all ranks, channels of ReluComposite / BootStrap dealt to the ranks")
    ap.add_argument("--batch", type=int, default=4, help="channels per batched ciphertext in ReluComposite / BootStrap")
    a = ap.parse_args()
    {"ours": run_ours, "layers": run_layers, "conv": run_conv, "full_vs_reference": run_full_vs_reference}[a.mode](a)
