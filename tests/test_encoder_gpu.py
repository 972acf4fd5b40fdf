"""GPU parity of the CKKS encoder / decoder against the reference's own PhantomCKKSEncoder
(oracle/_ref, built unmodified) and against the exact host decoder tests/ckks_host.py.

The encoder is floating point up to the rounding step: the special FFT is written with the reference's butterfly
expressions, so the residues are expected to be IDENTICAL; the stated tolerance, should a toolchain ever contract
the complex products differently, is one unit in the last place of the FFT output (|difference of an integer
coefficient| <= 2^-50 * |coefficient|), checked through the decoded values: |decode(encode(x)) - x| < 2^-40 * max|x|
at scale 2^50.
"""
import ctypes as C

import numpy as np
import pytest

import orc

pytestmark = pytest.mark.gpu


def _setup(fhe, logN, bits, nP):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    import refgpu
    if not refgpu.available():
        pytest.skip("reference GPU library not built (oracle/Makefile.ref)")
    ref = refgpu.Ref(logN, bits, nP)
    ctx = fhe.Context(logN, ref.primes, nP)
    return torch, refgpu, ref, ctx, fhe.Encoder(ctx)


def _ref_encode(refgpu, ref, v, scale, chain, mode, limbs):
    import torch
    out = torch.zeros((limbs, ref.N), dtype=torch.int64, device="cuda")
    f = np.ascontiguousarray(v.astype(np.complex128)).view(np.float64)
    n = refgpu.lib().rref_encode_raw(ref.h, f.ctypes.data_as(C.POINTER(C.c_double)), v.size, scale, chain, mode, out.data_ptr())
    assert n == limbs, refgpu.lib().rref_last_error()
    return out


@pytest.mark.parametrize("logN", [12, 14, 16])
def test_encode_matches_reference(fhe, logN):
    bits = [60, 50, 50, 50, 50, 60, 60]
    torch, refgpu, ref, ctx, enc = _setup(fhe, logN, bits, 2)
    Q = len(bits) - 2
    rng = np.random.default_rng(logN)
    slots = (1 << logN) // 2
    cases = [(rng.uniform(-1, 1, slots) + 1j * rng.uniform(-1, 1, slots), 2.0 ** 50, 1),      # full vector, top level
             (rng.uniform(-8, 8, slots // 4) + 0j, 2.0 ** 40, 3),                            # short vector: zero padded
             (rng.uniform(-1, 1, slots) + 1j * rng.uniform(-1, 1, slots), 2.0 ** 100, 1),     # 128-bit coefficients
             (rng.uniform(-1, 1, slots) + 0j, 2.0 ** 150, 1),                                # multi-word coefficients
             (np.array([1.5 - 2.25j]), 2.0 ** 30, 2)]                                        # a single value
    for v, scale, chain in cases:
        Ql = Q - chain + 1
        for ext in (False, True):
            ours = enc.encode(v, Ql, scale, ext=ext)
            theirs = _ref_encode(refgpu, ref, v, scale, chain, 1 if ext else 0, Ql + (2 if ext else 0))
            assert torch.equal(ours, theirs), "encode%s differs from the reference (scale 2^%d, chain %d)" % (
                "_ext" if ext else "", int(np.log2(scale)), chain)
    # error behaviour of encode_internal (src/ckks.cu:117-133,163-166)
    with pytest.raises(fhe.B200FheError):
        enc.encode(np.zeros(slots + 1, dtype=np.complex128), Q, 2.0 ** 40)
    with pytest.raises(fhe.B200FheError):
        enc.encode(np.ones(4, dtype=np.complex128), 1, 2.0 ** 70)            # scale out of bounds for one 60-bit prime
    with pytest.raises(fhe.B200FheError):
        enc.encode(np.full(slots, 1e30, dtype=np.complex128), 2, 2.0 ** 60)   # encoded values are too large


def test_encode_sparse_matches_reference(fhe):
    bits = [60, 50, 50, 50, 60]
    torch, refgpu, ref, ctx, enc = _setup(fhe, 13, bits, 1)
    Q = len(bits) - 1
    rng = np.random.default_rng(5)
    for n in (64, 256, 2048):   # (the reference's decompose launch has zero blocks below 2n * limbs = 128)
        v = rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)
        for ext in (False, True):
            ours = enc.encode(v, Q, 2.0 ** 45, ext=ext, sparse=True)
            theirs = _ref_encode(refgpu, ref, v, 2.0 ** 45, 1, 3 if ext else 2, Q + (1 if ext else 0))
            assert torch.equal(ours, theirs), "encode_sparse%s differs (n=%d)" % ("_ext" if ext else "", n)
            if not ext:  # only every (slots/n)-th coefficient is populated
                coef = ours.clone()
                ctx.ntt_inverse(coef, Q)
                mask = torch.ones(ctx.N, dtype=torch.bool, device="cuda")
                mask[::ctx.N // (2 * n)] = False
                assert int((coef[:, mask] != 0).sum().item()) == 0


def test_decode_round_trip_and_host_decoder(fhe):
    import ckks_host
    bits = [60, 50, 50, 60]
    torch, refgpu, ref, ctx, enc = _setup(fhe, 12, bits, 1)
    o = orc.Ctx(12, ref.primes, 1)
    rng = np.random.default_rng(9)
    slots = 2048
    x = rng.uniform(-4, 4, slots) + 1j * rng.uniform(-4, 4, slots)
    for Ql, scale in ((3, 2.0 ** 50), (2, 2.0 ** 40), (1, 2.0 ** 30), (3, 2.0 ** 100)):
        pt = enc.encode(x, Ql, scale)
        got = enc.decode(pt, Ql, scale)
        tol = max(2.0 ** -40 * np.abs(x).max(), 4.0 / scale * slots)     # rounding to integers: <= slots/2 ulp of 1/scale
        assert np.abs(got - x).max() < tol, (Ql, np.log2(scale), np.abs(got - x).max())
        host = ckks_host.decode(o, pt.cpu().numpy().view(np.uint64).reshape(Ql, ctx.N), scale)
        assert np.abs(got - host).max() < 2.0 ** -40 * np.abs(x).max() + 1e-12     # exact CRT on the host, numpy FFT
    # extended basis: the same values over Q_l and P
    pt = enc.encode(x, 2, 2.0 ** 45, ext=True)
    got = enc.decode(pt, 2, 2.0 ** 45, ext=True)
    assert np.abs(got - x).max() < 2.0 ** -30
    # against the reference decoder where it works on this toolchain: ONE limb (its multi-word CRT composition is
    # miscompiled by nvcc 12.9) -- the inverse NTT, the conversion to double, the forward special FFT
    # and the bit reversal are compared bit for bit
    pt = enc.encode(x, 1, 2.0 ** 30)
    theirs = np.empty(slots, dtype=np.complex128)
    rc = refgpu.lib().rref_decode_raw(ref.h, pt.data_ptr(), 3, 2.0 ** 30, theirs.view(np.float64).ctypes.data_as(C.POINTER(C.c_double)))
    assert rc == 0, refgpu.lib().rref_last_error()
    ours = enc.decode(pt, 1, 2.0 ** 30)
    assert np.array_equal(ours, theirs), np.abs(ours - theirs).max()


@pytest.mark.parametrize("n", [2, 64, 1024, 2048])
def test_encode_sparse_batch_equals_single_encodes(fhe, n):
    """the one-pass batch encoder (one CTA per message, whole FFT in shared memory) gives the residues of `batch`
    separate encode_sparse calls: 64-bit and 128-bit coefficients, masked-constant messages like DNN::Conv's weights"""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    logN = 13
    bits = [60, 59, 59, 59, 60, 60]
    primes = fhe.coeff_modulus_create(1 << logN, bits)
    ctx = fhe.Context(logN, primes, 2)
    enc = fhe.Encoder(ctx)
    rng = np.random.default_rng(n)
    for scale, Ql in ((2.0 ** 59, 4), (2.0 ** 100, 3)):
        vals = rng.uniform(-1, 1, (7, n)) + 0j
        vals[1] = 0.37 * (rng.uniform(0, 1, n) < 0.6)          # constant under a mask
        vals[2] = rng.uniform(-1, 1, n) + 1j * rng.uniform(-1, 1, n)
        vals[3] = 0.0
        got = enc.encode_sparse_batch(vals, Ql, scale)
        for z in range(vals.shape[0]):
            want = enc.encode(vals[z], Ql, scale, sparse=True)
            assert bool((got[z] == want).all()), "batch item %d differs (n=%d, scale=%g)" % (z, n, scale)
