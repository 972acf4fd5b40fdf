#!/usr/bin/env python
"""CKKS bootstrap on the reference's own configuration (bootstrapping_example.cu `simple`: N=2^16, |Q|=30
(60 + 29x59 bit), |P|=10 (60 bit), level budget {2,2}, 2^15 slots, scale 2^59): ours next to the reference's GPU
build on the same B200.

  --mode parity     keys and plaintext diagonals come from the reference library (oracle/_ref, zero-copy); checks
                    that every limb of our EvalBootstrap output equals the reference's, reports the decrypted
                    precision, and times both.
  --mode synthetic  uniform random keys / diagonals of the right shapes (no reference needed): timing only.
Prints one JSON object.  Used by bench.py (extra.bootstrap) and tests/ (TEST / BENCH infrastructure)."""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def ev_time(torch, fn, iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def host_e2e(torch, fhe, boot, t_in, Ql, scale, noise, Q, N, iters=3):
    """b200fhe_boot_eval_host: the input ciphertext in pinned host memory -> H2D, EvalBootstrap, D2H of the result"""
    L = fhe.lib()
    hin = torch.empty((2, Ql, N), dtype=torch.int64).pin_memory()
    hin.copy_(t_in.reshape(2, Ql, N).cpu())
    hout = torch.empty((2, Q, N), dtype=torch.int64).pin_memory()
    desc = fhe.CtDesc()

    def call():
        fhe._check(L.b200fhe_boot_eval_host(boot.h, hin.data_ptr(), Ql, scale, noise, hout.data_ptr(), hout.numel(), C.byref(desc)))
    call()
    t0 = time.perf_counter()
    for _ in range(iters):
        call()
    dt = (time.perf_counter() - t0) / iters
    return {"value": round(dt * 1e3, 3), "unit": "ms", "h2d_bytes_per_step": hin.numel() * 8,
            "d2h_bytes_per_step": int(2 * desc.size_Ql * N * 8), "levels_out": int(desc.size_Ql),
            "api": "b200fhe_boot_eval_host (pinned host ciphertext in and out; keys and diagonals device-resident)"}


def throughput_8(torch, fhe, ctx, boot, make_ct):
    """bootstrap batch = 8 per GPU (throughput): 8 independent ciphertexts on 2 / 4 lanes (one stream, pool and
    host thread each; keys and diagonals shared) against 8 bootstraps one after the other on one stream"""
    try:
        enc = fhe.Encoder(ctx)
        dnn = fhe.DNN(boot.ev, enc, 2.0 ** 59)
        cts = [make_ct() for _ in range(8)]
        t = fhe.TensorCT(cts, 0, 0, len(cts))   # width 0: a plain list of ciphertexts
        dnn.bootstrap(t, boot)
        c0 = time.process_time()
        serial = ev_time(torch, lambda: dnn.bootstrap(t, boot), 2)
        host_serial = (time.process_time() - c0) * 1e3 / 2
        res = {"ciphertexts": 8, "one_stream_ms": round(serial, 2), "one_stream_bootstraps_per_s": round(8e3 / serial, 2),
               "one_stream_host_cpu_ms_per_pass": round(host_serial, 1)}
        for lanes in (2, 4):
            dnn.set_lanes(boot, lanes)
            dnn.bootstrap(t, boot)
            c0 = time.process_time()
            ms = ev_time(torch, lambda: dnn.bootstrap(t, boot), 2)
            res["lanes%d_host_cpu_ms_per_pass" % lanes] = round((time.process_time() - c0) * 1e3 / 2, 1)   # all threads
            res["lanes%d_ms" % lanes] = round(ms, 2)
            res["lanes%d_bootstraps_per_s" % lanes] = round(8e3 / ms, 2)
        dnn.set_lanes(boot, 0)
        dnn.close()
        # the same 8 ciphertexts as batched handles: one launch set per step over the whole batch
        for bs in (4, 8):
            groups = [boot.ev.pack(cts[i:i + bs]) for i in range(0, 8, bs)]
            run = lambda: [boot.eval(g) for g in groups]   # noqa: E731
            run()
            ms = ev_time(torch, run, 2)
            res["batch%d_ms" % bs] = round(ms, 2)
            res["batch%d_bootstraps_per_s" % bs] = round(8e3 / ms, 2)
        return res
    except Exception as e:
        return {"error": repr(e)}


def chain_bits(levels_after, nP):
    depth = levels_after + 18
    return [60] + [59] * depth + [60] * nP


def run_parity(logN, levels_after, nP, iters, decode=True):
    import torch
    import phantom_fhe_boot_b200 as fhe
    import refgpu
    import orc
    R = refgpu.lib()
    N = 1 << logN
    bits = chain_bits(levels_after, nP)
    Q = len(bits) - nP
    out = {"mode": "parity", "config": "N=2^%d |Q|=%d |P|=%d budget {2,2} slots=%d scale=2^59" % (logN, Q, nP, N // 2)}
    t0 = time.time()
    ref = refgpu.Ref(logN, bits, nP)
    out["ref_context_s"] = round(time.time() - t0, 2)
    t0 = time.time()
    rb = refgpu.RefBoot(ref, (2, 2), 2.0 ** 59)
    out["ref_setup_keygen_s"] = round(time.time() - t0, 2)
    t0 = time.time()
    ctx = fhe.Context(logN, ref.primes, nP)
    boot = fhe.Bootstrapper(ctx, (2, 2))
    out["our_context_s"] = round(time.time() - t0, 2)
    rots = boot.rotation_indices()
    assert rots == rb.rotation_indices()
    n_diag = 0
    for which in (0, 1):
        for s, row in enumerate(rb.diagonals(which)):
            for i, (p, chain, scale, limbs) in enumerate(row):
                boot.set_diagonal(which, s, i, p, Q - (chain - 1), scale)
                n_diag += 1
    keys = rb.install_keys(boot.ev, rots)
    out["galois_keys"], out["diagonals"] = len(keys), n_diag
    free, total = torch.cuda.mem_get_info()
    out["hbm_used_GiB_after_setup"] = round((total - free) / 2 ** 30, 1)

    rng = np.random.default_rng(42)
    x = rng.uniform(-1.0, 1.0, N // 2)
    rct = R.rref_encrypt(ref.h, x.ctypes.data_as(C.POINTER(C.c_double)), len(x), 2.0 ** 59, 1)
    assert rct, R.rref_last_error()
    while R.rref_ct_limbs(rct) > 3 or R.rref_ct_noise(rct) != 2:      # bootstrapping_example.cu:150-153
        refgpu.ck(R.rref_boot_mult_const(rb.h, rct, 1.0))
    t_in = ref.ct_tensor(rct)
    ours_in = boot.ev.import_ct(t_in, R.rref_ct_limbs(rct), scale=R.rref_ct_scale(rct), noise=R.rref_ct_noise(rct))

    rout = rb.ct(R.rref_boot_eval, rct)
    ref.sync()
    n0 = fhe.launch_count()
    fhe.trace_reset()
    o = boot.eval(ours_in)
    torch.cuda.synchronize()
    out["our_launches_per_bootstrap"] = fhe.launch_count() - n0
    out["trace"] = fhe.trace_read()
    a, b = o.tensor(), ref.ct_tensor(rout)
    torch.cuda.synchronize()
    d = o.desc()
    out["levels_in"], out["levels_out"] = int(R.rref_ct_limbs(rct)), int(d.size_Ql)
    out["bit_exact_vs_reference"] = bool(a.shape == b.shape and (a == b).all().item()) and d.scale == R.rref_ct_scale(rout)
    del a, b
    if decode:
        got = ref.decrypt_slots(rout, orc.Ctx(logN, ref.primes, nP))
        err = np.abs(got.real - x)
        out["max_abs_err"] = float(err.max())
        out["precision_bits_mean"] = float(-np.log2(err.mean()))

    # timing: warm both once more, then CUDA events on each library's own stream
    f = C.c_float()
    refgpu.ck(R.rref_time_bootstrap(rb.h, rct, 1, C.byref(f)))
    refgpu.ck(R.rref_time_bootstrap(rb.h, rct, iters, C.byref(f)))
    out["reference_gpu_ms"] = round(f.value, 2)
    boot.eval(ours_in)
    ms = ev_time(torch, lambda: boot.eval(ours_in), iters)
    out["ours_ms"] = round(ms, 2)
    t0 = time.perf_counter()
    boot.eval(ours_in)
    torch.cuda.synchronize()
    out["ours_wall_ms"] = round((time.perf_counter() - t0) * 1e3, 2)
    out["speedup_vs_reference_gpu"] = round(f.value / ms, 3)
    try:   # the same bootstrap replayed as a CUDA graph (b200fhe_boot_graph_capture): bit-exactness and latency
        g = boot.capture(boot.ev.import_ct(t_in, R.rref_ct_limbs(rct), scale=R.rref_ct_scale(rct), noise=R.rref_ct_noise(rct)))
        got = g.launch().tensor()
        torch.cuda.synchronize()
        want = ref.ct_tensor(rout)
        out["cuda_graph"] = {"bit_exact_vs_reference": bool(got.shape == want.shape and (got == want).all().item()),
                             "ms": round(ev_time(torch, lambda: g.launch(), iters), 2)}
        del got, want
        g.close()
    except Exception as e:
        out["cuda_graph"] = {"error": repr(e)}
    out["e2e"] = host_e2e(torch, fhe, boot, t_in, int(R.rref_ct_limbs(rct)), R.rref_ct_scale(rct), int(R.rref_ct_noise(rct)), Q, N)
    out["throughput_8_per_gpu"] = throughput_8(torch, fhe, ctx, boot, lambda: boot.ev.import_ct(
        t_in, R.rref_ct_limbs(rct), scale=R.rref_ct_scale(rct), noise=R.rref_ct_noise(rct)))
    out["our_pool_GiB"] = round(boot.ev.pool_bytes() / 2 ** 30, 2)
    # give the ~55 GiB of reference keys / diagonals back before returning
    del o, ours_in
    boot.close()
    ctx.close()
    for c in (rout, rct):
        R.rref_ct_free(c)
    R.rref_boot_destroy(rb.h)
    R.rref_destroy(ref.h)
    torch.cuda.empty_cache()
    return out


def synthetic_setup(logN, levels_after, nP, device=0, slots=0, seed=None):
    """context, bootstrapper with uniform random keys / diagonals of the right shapes, and an input ciphertext;
    seed != None: the same values on every process that passes the same seed (ranks of a limb-sharded bootstrap)"""
    import torch
    import phantom_fhe_boot_b200 as fhe
    N = 1 << logN
    bits = chain_bits(levels_after, nP)
    Q = len(bits) - nP
    primes = fhe.coeff_modulus_create(N, bits)
    ctx = fhe.Context(logN, primes, nP, device=device)
    boot = fhe.Bootstrapper(ctx, (2, 2), num_slots=slots)
    gen = None
    if seed is not None:
        gen = torch.Generator(device="cuda")
        gen.manual_seed(seed)

    def rnd(*shape):
        return torch.randint(0, 1 << 58, shape, dtype=torch.int64, device="cuda", generator=gen)
    beta = ctx.beta(Q)
    sf = boot.ev.scaling_factors(0)
    for which in (0, 1):
        n_stages = boot.stage_info(which, 0)[0]
        for s in range(n_stages):
            _, n, Ql = boot.stage_info(which, s)
            block = rnd(n, Ql + nP, N)
            for i in range(n):
                boot.set_diagonal(which, s, i, block[i], Ql, sf[Q - Ql])
    boot.ev.set_relin_key([rnd(2, Q + nP, N) for _ in range(beta)])
    ks = sorted({fhe.auto_index(logN, r) for r in boot.rotation_indices()} | {2 * N - 1})
    for k in ks:
        boot.ev.set_galois_key(k, [rnd(2, Q + nP, N) for _ in range(beta)])
    ct = boot.ev.import_ct(rnd(2, 3, N), 3, scale=sf[Q - 3] ** 2, noise=2)
    return ctx, boot, ct, {"Q": Q, "N": N, "sf": sf, "galois_keys": len(ks), "rnd": rnd}


def run_synthetic(logN, levels_after, nP, iters, device=0, profile=False, slots=0, throughput=False):
    import torch
    import phantom_fhe_boot_b200 as fhe
    ctx, boot, ct, info = synthetic_setup(logN, levels_after, nP, device, slots)
    N, Q, sf, rnd = info["N"], info["Q"], info["sf"], info["rnd"]
    out = {"mode": "synthetic", "config": "N=2^%d |Q|=%d |P|=%d budget {2,2} slots=%d; uniform random keys/diagonals" % (logN, Q, nP, slots or N // 2)}
    out["galois_keys"] = info["galois_keys"]
    n0 = fhe.launch_count()
    fhe.trace_reset()
    boot.eval(ct)
    torch.cuda.synchronize()
    out["our_launches_per_bootstrap"] = fhe.launch_count() - n0
    out["trace"] = fhe.trace_read()
    boot.eval(ct)
    torch.cuda.synchronize()
    if profile:  # ncu --profile-from-start off: exactly one bootstrap inside the profiled range
        torch.cuda.profiler.start()
        boot.eval(ct)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    out["ours_ms"] = round(ev_time(torch, lambda: boot.eval(ct), iters), 2)
    # host time to ISSUE one bootstrap (the call returns before the GPU finishes): launch-bound when close to ours_ms
    torch.cuda.synchronize()
    t0 = time.time()
    boot.eval(ct)
    out["host_issue_ms"] = round((time.time() - t0) * 1e3, 2)
    torch.cuda.synchronize()
    out["our_pool_GiB"] = round(boot.ev.pool_bytes() / 2 ** 30, 2)
    free, total = torch.cuda.mem_get_info()
    out["hbm_used_GiB"] = round((total - free) / 2 ** 30, 1)
    try:
        out["e2e"] = host_e2e(torch, fhe, boot, ct.tensor(), 3, sf[Q - 3] ** 2, 2, Q, N)
    except Exception as e:
        out["e2e"] = {"error": repr(e)}
    if throughput:
        out["throughput_8_per_gpu"] = throughput_8(torch, fhe, ctx, boot, lambda: boot.ev.import_ct(rnd(2, 3, N), 3, scale=sf[Q - 3] ** 2, noise=2))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="parity", choices=["parity", "synthetic"])
    ap.add_argument("--logN", type=int, default=16)
    ap.add_argument("--levels-after", type=int, default=11)
    ap.add_argument("--nP", type=int, default=10)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--no-decode", action="store_true")
    ap.add_argument("--slots", type=int, default=0, help="synthetic mode: sparse packing (0 = N/2)")
    ap.add_argument("--throughput", action="store_true", help="synthetic mode: also 8 ciphertexts per GPU on lanes")
    ap.add_argument("--profile", action="store_true", help="synthetic mode: cudaProfilerStart/Stop around one bootstrap")
    a = ap.parse_args()
    if a.mode == "parity":
        r = run_parity(a.logN, a.levels_after, a.nP, a.iters, not a.no_decode)
    else:
        r = run_synthetic(a.logN, a.levels_after, a.nP, a.iters, profile=a.profile, slots=a.slots, throughput=a.throughput)
    print(json.dumps(r))


if __name__ == "__main__":
    main()
