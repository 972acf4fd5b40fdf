#!/usr/bin/env python
"""Inverse transform at N = 2^16 over few limbs: the two-pass kernels against the one-kernel cluster transform
(ntt_inv_cluster, b200fhe_tune_cluster_ntt), bit-exactness of one against the other and warm timings per launch size;
then HMult+relin, rotate and the synthetic bootstrap with the cluster transform off / on.  Prints one JSON object."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    import torch
    import phantom_fhe_boot_b200 as fhe
    L = fhe.lib()
    logN, Q, P = 16, 30, 10
    N = 1 << logN
    primes = fhe.coeff_modulus_create(N, [60] + [59] * (Q - 1) + [60] * P)
    ctx = fhe.Context(logN, primes, P)

    def timed(fn, iters=50):
        for _ in range(5):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return round(e0.elapsed_time(e1) / iters * 1e3, 2)

    out = {"unit": "us", "config": "N=2^16, inverse transform of `limbs` limbs x `polys` polynomials, warm, back to back", "sizes": {}}
    ok = True
    for limbs, polys in ((1, 1), (1, 2), (2, 2), (4, 1), (6, 1), (8, 1), (10, 1), (12, 1), (16, 1), (10, 2), (8, 2), (30, 1)):
        x0 = torch.randint(0, 1 << 58, (polys, limbs, N), dtype=torch.int64, device="cuda")
        res = {}
        outs = {}
        for name, knob in (("two_pass", 0), ("cluster", 64)):
            L.b200fhe_tune_cluster_ntt(knob)
            x = x0.clone()
            ctx.ntt_inverse(x, limbs, npoly=polys, poly_stride=limbs * N)
            torch.cuda.synchronize()
            outs[name] = x
            y = x0.clone()
            res[name] = timed(lambda: ctx.ntt_inverse(y, limbs, npoly=polys, poly_stride=limbs * N))
        same = bool((outs["two_pass"] == outs["cluster"]).all())
        ok = ok and same
        res["bit_exact"] = same
        res["speedup"] = round(res["two_pass"] / res["cluster"], 2)
        out["sizes"]["%dx%d" % (limbs, polys)] = res
    out["all_bit_exact"] = ok

    def rnd(*shape):
        return torch.randint(0, 1 << 58, shape, dtype=torch.int64, device="cuda")
    Ql = Q
    beta = ctx.beta(Ql)
    evk = [rnd(2, Q + P, N) for _ in range(beta)]
    a_, b_, o_ = rnd(2, Ql, N), rnd(2, Ql, N), rnd(2, Ql, N)
    ct = rnd(2, Ql, N)
    comp = {}
    ref_out = None
    for name, knob in (("two_pass", 0), ("cluster_le_16", 16), ("cluster_le_10", 10)):
        L.b200fhe_tune_cluster_ntt(knob)
        ctx.hmult_relin(Ql, o_, a_, b_, evk)
        torch.cuda.synchronize()
        if ref_out is None:
            ref_out = o_.clone()
        comp[name] = {"hmult_relin_us": timed(lambda: ctx.hmult_relin(Ql, o_, a_, b_, evk), 100),
                      "hmult_bit_exact": bool((o_ == ref_out).all())}
    out["hmult_relin"] = comp
    ctx.close()
    del evk, a_, b_, o_, ct
    torch.cuda.empty_cache()
    import boot_bench
    _, boot, bct, info = boot_bench.synthetic_setup(16, 11, 10, seed=7)
    bres = {}
    want = None
    for name, knob in (("two_pass", 0), ("cluster_le_16", 16), ("cluster_le_10", 10)):
        L.b200fhe_tune_cluster_ntt(knob)
        got = boot.eval(bct).tensor().clone()
        torch.cuda.synchronize()
        if want is None:
            want = got
        bres[name] = {"bootstrap_ms": round(boot_bench.ev_time(torch, lambda: boot.eval(bct), 5), 3), "bit_exact": bool((got == want).all())}
    out["bootstrap"] = bres
    L.b200fhe_tune_cluster_ntt(-1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
