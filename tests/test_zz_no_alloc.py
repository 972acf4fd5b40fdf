"""GPU: b200fhe_rescale_noalloc (static __device__ scratch, allocates nothing) against b200fhe_rescale, through the
C ABI, on two calls in a row.  Integer work: the two entry points must agree bit for bit (tolerance 0, as in the other
parity tests); the oracle is the second checker.  Collected last on purpose."""
import numpy as np
import pytest

import orc


def _eq(a, b):
    return np.array_equal(np.asarray(a).reshape(-1), np.asarray(b).reshape(-1))


@pytest.mark.gpu
def test_rescale_noalloc_matches_rescale_two_calls_in_a_row(fhe):
    import torch
    assert torch.cuda.is_available()
    L = fhe.lib()
    for logN, bits, nP in ((12, [60, 59, 59, 59, 60, 60], 2), (16, [60, 59, 59, 60], 1)):
        N = 1 << logN
        primes = orc.coeff_modulus_create(N, bits)
        ctx, o = fhe.Context(logN, primes, nP, device=0), orc.Ctx(logN, primes, nP)
        nQ = len(primes) - nP
        rng = np.random.default_rng(20 + logN)
        # two calls in a row on different inputs and levels: the second finds the first one's data in the scratch
        for call, (Lq, polys) in enumerate(((nQ, 2), (nQ - 1 if nQ > 2 else nQ, 2), (nQ, 1))):
            src = orc.rand_residues(rng, primes[:Lq], N, lead=(polys,))
            dsrc = fhe.to_dev(src)
            want = fhe.to_dev(np.zeros((polys, Lq - 1, N), dtype=np.uint64))
            got = fhe.to_dev(np.full((polys, Lq - 1, N), 0xDEADBEEF, dtype=np.uint64))
            ctx.rescale(Lq, want, dsrc, polys)
            rc = L.b200fhe_rescale_noalloc(ctx.h, Lq, got.data_ptr(), dsrc.data_ptr(), polys, fhe._stream())
            assert rc == 0, L.b200fhe_last_error().decode()
            torch.cuda.synchronize()
            assert _eq(fhe.to_host(got), fhe.to_host(want)), "no-alloc rescale differs from b200fhe_rescale (call %d)" % call
            assert _eq(fhe.to_host(dsrc), src), "source clobbered"
            if logN == 12:
                assert _eq(fhe.to_host(got), o.rescale(Lq, src)), "no-alloc rescale differs from the oracle"
        # beyond the static maximum: a clear error, nothing launched
        big = 3 if logN == 16 else (131072 // N) + 1
        dummy = fhe.to_dev(np.zeros(8, dtype=np.uint64))
        rc = L.b200fhe_rescale_noalloc(ctx.h, nQ, dummy.data_ptr(), dsrc.data_ptr(), big, fhe._stream())
        assert rc == -1 and b"static scratch" in L.b200fhe_last_error()
