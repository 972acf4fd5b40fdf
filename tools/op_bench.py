#!/usr/bin/env python
"""Warm (back-to-back, L2 in its natural state) device timings of the composite operations at the shapes the
bootstrap uses them, through the C ABI.  The ncu launch lists under profiles/ are cold-cache and serialised; this
is the other view: what one operation costs inside a stream of them.  Prints one JSON object."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--logN", type=int, default=16)
    ap.add_argument("--iters", type=int, default=50)
    a = ap.parse_args()
    import torch
    import phantom_fhe_boot_b200 as fhe
    N = 1 << a.logN
    Q, P = 30, 10
    primes = fhe.coeff_modulus_create(N, [60] + [59] * (Q - 1) + [60] * P)
    ctx = fhe.Context(a.logN, primes, P)

    def rnd(*shape):
        return torch.randint(0, 1 << 58, shape, dtype=torch.int64, device="cuda")

    def timed(fn, iters=a.iters):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return round(e0.elapsed_time(e1) / iters * 1e3, 2)   # us

    out = {"unit": "us", "config": "N=2^%d |Q|=%d |P|=%d" % (a.logN, Q, P)}
    for Ql in (30, 24, 16):
        beta = ctx.beta(Ql)
        QlP = Ql + P
        r = {}
        for polys in (2, 4):
            src, dst = rnd(polys, Ql, N), rnd(polys, Ql - 1, N)
            r["rescale_polys%d" % polys] = timed(lambda: ctx.rescale(Ql, dst, src, polys=polys))
        x = rnd(4, Ql, N)
        r["ntt_fwd_4polys"] = timed(lambda: ctx.ntt_forward(x, Ql, npoly=4, poly_stride=Ql * N))
        r["ntt_inv_4polys"] = timed(lambda: ctx.ntt_inverse(x, Ql, npoly=4, poly_stride=Ql * N))
        c2, up, cx, ct = rnd(Ql, N), rnd(beta, QlP, N), rnd(2, QlP, N), rnd(2, Ql, N)
        evk = [rnd(2, Q + P, N) for _ in range(beta)]
        r["modup"] = timed(lambda: ctx.modup(Ql, up, c2))
        r["inner_prod"] = timed(lambda: ctx.inner_prod(Ql, cx, up, evk))
        r["moddown_2polys"] = timed(lambda: ctx.moddown_from_ntt(Ql, ct, cx))
        r["keyswitch"] = timed(lambda: ctx.keyswitch(Ql, ct, c2, evk))
        a_, b_, o_ = rnd(2, Ql, N), rnd(2, Ql, N), rnd(2, Ql, N)
        r["hmult_relin"] = timed(lambda: ctx.hmult_relin(Ql, o_, a_, b_, evk))
        out["Ql=%d" % Ql] = r
    print(json.dumps(out))


if __name__ == "__main__":
    main()
