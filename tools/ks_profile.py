#!/usr/bin/env python
"""One HMult+relin and one EvalRotateFused-shaped key switch at config K between cudaProfilerStart/Stop, for
`ncu --profile-from-start off` launch lists and `--set full` captures of the key-switch kernels (profiles/r02_*)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import phantom_fhe_boot_b200 as fhe
    logN, Q, P = 16, 30, 10
    Ql = int(sys.argv[1]) if len(sys.argv) > 1 else Q
    N = 1 << logN
    primes = fhe.coeff_modulus_create(N, [60] + [59] * (Q - 1) + [60] * P)
    ctx = fhe.Context(logN, primes, P)

    def rnd(*shape):
        return torch.randint(0, 1 << 58, shape, dtype=torch.int64, device="cuda")
    beta = ctx.beta(Ql)
    evk = [rnd(2, Q + P, N) for _ in range(beta)]
    a_, b_, o_ = rnd(2, Ql, N), rnd(2, Ql, N), rnd(2, Ql, N)
    for _ in range(3):
        ctx.hmult_relin(Ql, o_, a_, b_, evk)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    ctx.hmult_relin(Ql, o_, a_, b_, evk)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()


if __name__ == "__main__":
    main()
