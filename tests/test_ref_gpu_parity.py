"""GPU parity against the REFERENCE ITSELF: the unmodified reference kernels / evaluator (oracle/_ref/libphantom_ref.so)
and libb200fhe.so run on identical device buffers, keys and ciphertexts on the B200; outputs must be bit-identical.
This pins the oracle and the product to "outputs of the reference itself run here" (the reference has no golden
vectors for this path)."""
import ctypes as C

import numpy as np
import pytest

import orc
import refgpu

pytestmark = pytest.mark.gpu


def _need():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    if not refgpu.available():
        pytest.skip("oracle/_ref/libphantom_ref.so not built (reference tree absent at build time)")
    return torch


def eq(a, b):
    return np.array_equal(np.asarray(a).reshape(-1), np.asarray(b).reshape(-1))


def dp(v):
    return v.ctypes.data_as(C.POINTER(C.c_double))


@pytest.mark.parametrize("logN,bits,nP", [(12, [60, 60, 60, 60], 1), (14, [60, 59, 59, 59, 60, 60], 2),
                                          (16, [60] + [59] * 5 + [60] * 2, 2), (15, [60] + [50] * 7 + [60], 1),
                                          (16, [60] + [59] * 35 + [60], 1),    # 36 limbs: the bench shape
                                          # the rest of benchmark/ntt_bench's sweep (N = 2^14..2^16 x L = 8, 16, 24, 36)
                                          (14, [60] + [59] * 7 + [60], 1), (14, [60] + [59] * 15 + [60], 1),
                                          (14, [60] + [59] * 23 + [60], 1), (14, [60] + [59] * 35 + [60], 1),
                                          (15, [60] + [59] * 7 + [60], 1), (15, [60] + [59] * 15 + [60], 1),
                                          (15, [60] + [59] * 23 + [60], 1), (15, [60] + [59] * 35 + [60], 1),
                                          (16, [60] + [59] * 7 + [60], 1), (16, [60] + [59] * 15 + [60], 1),
                                          (16, [60] + [59] * 23 + [60], 1)])
def test_ntt_vs_reference_kernels(fhe, logN, bits, nP):
    torch = _need()
    ref = refgpu.Ref(logN, bits, nP)
    ctx = fhe.Context(logN, ref.primes, nP)
    N, nQP = 1 << logN, len(bits)
    rng = np.random.default_rng(21)
    x = orc.rand_residues(rng, ref.primes, N)
    ours, theirs = fhe.to_dev(x), fhe.to_dev(x)
    ctx.ntt_forward(ours, nQP)
    refgpu.ck(refgpu.lib().rref_ntt_forward(ref.h, theirs.data_ptr(), nQP, 0))
    ref.sync()
    torch.cuda.synchronize()
    assert eq(fhe.to_host(ours), fhe.to_host(theirs)), "forward NTT differs from the reference kernel"
    ours, theirs = fhe.to_dev(x), fhe.to_dev(x)
    ctx.ntt_inverse(ours, nQP)
    refgpu.ck(refgpu.lib().rref_ntt_inverse(ref.h, theirs.data_ptr(), nQP, 0))
    ref.sync()
    torch.cuda.synchronize()
    assert eq(fhe.to_host(ours), fhe.to_host(theirs)), "inverse NTT differs from the reference kernel"
    o = orc.Ctx(logN, ref.primes, nP)  # the oracle agrees too: pins the oracle to the reference's kernels
    assert eq(fhe.to_host(theirs), o.ntt_inv(x))


@pytest.mark.parametrize("logN,bits,nP,drop", [(12, [60, 59, 59, 59, 59, 60, 60], 2, [0, 1, 2]),
                                               (13, [60, 50, 50, 50, 50, 50, 50, 60, 60, 60], 3, [0, 3]),
                                               # BASELINE configs[2] at full size: K' (L = 36, alpha = 12, dnum = 3) and
                                               # the reference's own benchmark/keyswitch_bench.cu shape (alpha = 15, dnum = 2)
                                               (16, [60] + [59] * 35 + [60] * 12, 12, [0, 13]),
                                               # config K itself: the bootstrap chain (|Q| = 30, alpha = 10, dnum = 3)
                                               (16, [60] + [59] * 29 + [60] * 10, 10, [0, 11]),
                                               (15, [60] + [50] * 29 + [60] * 15, 15, [0])])
def test_keyswitch_stages_vs_reference(fhe, logN, bits, nP, drop):
    torch = _need()
    ref = refgpu.Ref(logN, bits, nP)
    L = refgpu.lib()
    ctx = fhe.Context(logN, ref.primes, nP)
    o = orc.Ctx(logN, ref.primes, nP)
    N, nQP = 1 << logN, len(bits)
    Q = nQP - nP
    rlk = L.rref_gen_relinkey(ref.h)
    assert rlk, L.rref_last_error()
    digits = ref.relinkey_digits(rlk)
    rng = np.random.default_rng(22)
    for dr in drop:
        Ql = Q - dr
        QlP = Ql + nP
        beta = (Ql + nP - 1) // nP
        if beta > len(digits):
            continue
        c2 = orc.rand_residues(rng, ref.primes[:Ql], N)
        up_o = fhe.to_dev(np.zeros((beta, QlP, N), dtype=np.uint64))
        up_r = fhe.to_dev(np.zeros((beta, QlP, N), dtype=np.uint64))
        dc2 = fhe.to_dev(c2)
        ctx.modup(Ql, up_o, dc2)
        refgpu.ck(L.rref_modup(ref.h, dr, up_r.data_ptr(), dc2.data_ptr()))
        ref.sync()
        torch.cuda.synchronize()
        assert eq(fhe.to_host(up_o), fhe.to_host(up_r)), "modup differs from the reference (drop=%d)" % dr
        assert eq(fhe.to_host(up_r), o.modup(Ql, c2)), "oracle modup differs from the reference"
        cx_o = fhe.to_dev(np.zeros((2, QlP, N), dtype=np.uint64))
        cx_r = fhe.to_dev(np.zeros((2, QlP, N), dtype=np.uint64))
        ctx.inner_prod(Ql, cx_o, up_o, digits[:beta])
        refgpu.ck(L.rref_inner_prod(ref.h, dr, cx_r.data_ptr(), up_r.data_ptr(), rlk))
        ref.sync()
        torch.cuda.synchronize()
        assert eq(fhe.to_host(cx_o), fhe.to_host(cx_r)), "inner product differs from the reference"
        ct_o = fhe.to_dev(np.zeros((Ql, N), dtype=np.uint64))
        ct_r = fhe.to_dev(np.zeros((Ql, N), dtype=np.uint64))
        cxo2, cxr2 = cx_o[0].clone(), cx_r[0].clone()
        ctx.moddown_from_ntt(Ql, ct_o, cxo2)
        refgpu.ck(L.rref_moddown_from_ntt(ref.h, dr, ct_r.data_ptr(), cxr2.data_ptr()))
        ref.sync()
        torch.cuda.synchronize()
        assert eq(fhe.to_host(ct_o), fhe.to_host(ct_r)), "moddown differs from the reference"
    L.rref_free_relinkey(rlk)


def test_evaluator_vs_reference_same_keys(fhe):
    """multiply_and_relin_inplace, rescale_to_next, EvalRotateFused on ciphertexts ENCRYPTED BY THE REFERENCE with
    keys GENERATED BY THE REFERENCE; every output limb must match.  Decrypted values: max-abs-error < 2^-20."""
    torch = _need()
    logN, bits, nP = 13, [60, 50, 50, 50, 50, 50, 60, 60], 2
    ref = refgpu.Ref(logN, bits, nP)
    L = refgpu.lib()
    ctx = fhe.Context(logN, ref.primes, nP)
    N, Q = 1 << logN, len(bits) - nP
    rlk = L.rref_gen_relinkey(ref.h)
    assert rlk, L.rref_last_error()
    rdig = ref.relinkey_digits(rlk)
    rng = np.random.default_rng(23)
    va, vb = rng.uniform(-1, 1, N // 2), rng.uniform(-1, 1, N // 2)
    scale = 2.0 ** 50
    a = L.rref_encrypt(ref.h, dp(va), N // 2, scale, 1)
    b = L.rref_encrypt(ref.h, dp(vb), N // 2, scale, 1)
    assert a and b, L.rref_last_error()
    ta, tb = ref.ct_tensor(a), ref.ct_tensor(b)
    assert tuple(ta.shape) == (2, Q, N)
    out = torch.zeros_like(ta)
    ctx.hmult_relin(Q, out, ta, tb, rdig)
    a2 = L.rref_ct_clone(a)
    refgpu.ck(L.rref_multiply_and_relin(ref.h, a2, b, rlk))
    tr = ref.ct_tensor(a2)
    assert tuple(tr.shape) == (2, Q, N)
    assert eq(fhe.to_host(out), fhe.to_host(tr)), "multiply_and_relin_inplace differs from the reference"
    refgpu.ck(L.rref_rescale(ref.h, a2))
    o = orc.Ctx(logN, ref.primes, nP)
    assert np.abs(ref.decrypt_slots(a, o) - va).max() < 2.0 ** -30   # fresh ciphertext (checks the host decoder)
    assert np.abs(ref.decrypt_slots(a2, o) - va * vb).max() < 2.0 ** -20
    res = torch.zeros((2, Q - 1, N), dtype=torch.int64, device="cuda")
    ctx.rescale(Q, res, out, 2)
    assert eq(fhe.to_host(res), fhe.to_host(ref.ct_tensor(a2))), "rescale_to_next differs from the reference"
    idx = np.array([1, 5, -3], dtype=np.int32)
    keys = L.rref_gen_rotkeys(ref.h, idx.ctypes.data_as(C.POINTER(C.c_int32)), len(idx))
    assert keys, L.rref_last_error()
    for i in idx:
        k = L.rref_auto_index(ref.h, int(i))
        key = L.rref_rotkey_get(keys, k)
        assert key, L.rref_last_error()
        kd = ref.relinkey_digits(key)
        rot = torch.zeros_like(tb)
        ctx.rotate_fused(Q, rot, tb, k, kd)
        outc = L.rref_ct_clone(b)
        refgpu.ck(L.rref_rotate_fused(ref.h, keys, b, outc, int(i)))
        assert eq(fhe.to_host(rot), fhe.to_host(ref.ct_tensor(outc))), "EvalRotateFused differs (index %d)" % i
        assert np.abs(ref.decrypt_slots(outc, o) - np.roll(vb, -int(i))).max() < 2.0 ** -20
        L.rref_ct_free(outc)
