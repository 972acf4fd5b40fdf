"""Serialization in the reference's byte formats: records written by libb200fhe.so load with the
reference's own load() and the other way round; saving the same object on both sides gives IDENTICAL bytes."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LOGN, BITS, NP = 12, [60, 50, 50, 50, 60, 60], 2


def _setup(fhe, need_ref=True):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    import refgpu
    if need_ref and not refgpu.available():
        pytest.skip("oracle/_ref/libphantom_ref.so not built")
    ref = refgpu.Ref(LOGN, BITS, NP)
    return torch, refgpu, ref, fhe.Context(LOGN, ref.primes, NP)


def _ref_save(refgpu, kind, obj):
    L = refgpu.lib()
    n = L.rref_save(kind, obj, None, 0)
    assert n > 0, L.rref_last_error()
    buf = (C.c_uint8 * n)()
    assert L.rref_save(kind, obj, buf, n) == n
    return bytes(buf)


def _ref_load(refgpu, kind, h, data):
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    o = refgpu.lib().rref_load(kind, h, buf, len(data))
    assert o, refgpu.lib().rref_last_error()
    return o


def test_formats_interchange_with_reference(fhe):
    torch, refgpu, ref, ctx = _setup(fhe)
    L = refgpu.lib()
    N, Q, QP = ctx.N, len(BITS) - NP, len(BITS)
    rng = np.random.default_rng(2)
    v = rng.uniform(-1, 1, N // 2)
    scale = 2.0 ** 40
    # ---- ciphertext: theirs -> bytes -> ours; ours -> bytes identical to theirs; ours -> bytes -> theirs
    a = L.rref_encrypt(ref.h, v.ctypes.data_as(C.POINTER(C.c_double)), N // 2, scale, 2)
    assert a, L.rref_last_error()
    theirs = _ref_save(refgpu, 0, a)
    h, t, used = fhe.load_ciphertext(theirs)
    assert used == len(theirs) == 58 + 2 * (Q - 1) * N * 8
    assert (h.chain_index, h.polys, h.N, h.limbs, h.scale, h.is_ntt_form) == (2, 2, N, Q - 1, scale, 1)
    assert torch.equal(t, ref.ct_tensor(a))
    ours = fhe.save_ciphertext(ctx, t, Q - 1, scale, noise=h.noise_scale_deg)
    assert ours == theirs, "ciphertext record differs from the reference's bytes"
    b = _ref_load(refgpu, 0, ref.h, ours)
    assert L.rref_ct_chain_index(b) == 2 and L.rref_ct_limbs(b) == Q - 1 and L.rref_ct_scale(b) == scale
    assert torch.equal(ref.ct_tensor(b), t)
    import orc
    o = orc.Ctx(LOGN, ref.primes, NP)     # (host decoder: the reference's own multi-limb decode is miscompiled)
    assert np.abs(ref.decrypt_slots(b, o).real - v).max() < 2.0 ** -20
    # ---- relinearisation key: both directions, identical bytes
    rlk = L.rref_gen_relinkey(ref.h)
    theirs = _ref_save(refgpu, 1, rlk)
    digits, used = fhe.load_eval_key(ctx, theirs)
    assert used == len(theirs) and len(digits) == Q // NP
    rd = ref.relinkey_digits(rlk)
    for d, p in zip(digits, rd):
        assert torch.equal(d.reshape(-1), refgpu.as_tensor(p if isinstance(p, int) else p.data_ptr(), 2 * QP * N))
    assert fhe.save_eval_key(ctx, digits) == theirs, "evaluation-key record differs from the reference's bytes"
    kg = fhe.KeyGenerator(ctx, bytes(range(32)))
    sk = torch.empty((QP, N), dtype=torch.int64, device="cuda")
    refgpu.ck(L.rref_secret_key(ref.h, sk.data_ptr()))
    kg.set_secret(sk)
    mine = kg.relin_key()
    torch.cuda.synchronize()
    rk = _ref_load(refgpu, 1, ref.h, fhe.save_eval_key(ctx, mine.digits))      # our key through their load()
    b2 = L.rref_ct_clone(a)
    refgpu.ck(L.rref_multiply_and_relin(ref.h, b2, a, rk))
    assert np.abs(ref.decrypt_slots(b2, o).real - v * v).max() < 2.0 ** -20
    # ---- secret key: theirs -> ours, identical bytes back, ours -> theirs
    theirs = _ref_save(refgpu, 2, ref.h)
    s = fhe.load_secret_key(ctx, theirs)
    assert torch.equal(s, sk)
    ours = fhe.save_secret_key(ctx, s)
    powers = int.from_bytes(theirs[:8], "little")       # the reference also stores s^2.. once it has computed them
    assert len(theirs) == 24 + powers * QP * N * 8 and theirs[8:24] == ours[8:24] and theirs[24:len(ours)] == ours[24:]
    if powers == 1:
        assert ours == theirs
    kg2 = fhe.KeyGenerator(ctx, bytes(range(1, 33)))
    _ref_load(refgpu, 2, ref.h, fhe.save_secret_key(ctx, kg2.secret()))         # the reference now holds OUR secret
    enc = fhe.Encoder(ctx)
    ct = kg2.encrypt_symmetric(enc.encode(v + 0j, Q, scale), Q)
    torch.cuda.synchronize()
    c = L.rref_ct_from_raw(ref.h, 1, 2, ct.data_ptr(), scale)
    assert np.abs(ref.decrypt_slots(c, o).real - v).max() < 2.0 ** -25


def test_round_trips_and_errors(fhe):
    torch, refgpu, ref, ctx = _setup(fhe, need_ref=False)
    N, Q, QP = ctx.N, len(BITS) - NP, len(BITS)
    kg = fhe.KeyGenerator(ctx, bytes(range(32)))
    enc = fhe.Encoder(ctx)
    x = np.linspace(-1, 1, N // 2) + 0j
    pt = enc.encode(x, Q - 1, 2.0 ** 45)
    data = fhe.save_plaintext(ctx, pt, Q - 1, 2.0 ** 45)
    ci, scale, t, used = fhe.load_plaintext(data)
    assert (ci, scale, used) == (2, 2.0 ** 45, len(data)) and torch.equal(t, pt)
    keys = kg.rotation_keys([1, 2, -5])
    blob = fhe.save_galois_keys(ctx, keys)
    back = fhe.load_galois_keys(ctx, blob)
    assert sorted(back) == sorted(keys)
    for k in keys:
        for d, p in zip(back[k], keys[k].digits):
            assert torch.equal(d.reshape(-1), refgpu.as_tensor(p, 2 * QP * N))
    ct = kg.encrypt_symmetric(pt, Q - 1)
    rec = fhe.save_ciphertext(ctx, ct, Q - 1, 2.0 ** 45)
    # truncated / corrupt / foreign records are refused, never read past the buffer
    for bad in (rec[:40], rec[:-8], b"\xff" * 58 + rec[58:], b""):
        with pytest.raises((fhe.B200FheError, ValueError)):
            fhe.load_ciphertext(bad)
    with pytest.raises(fhe.B200FheError):
        fhe.load_galois_keys(ctx, b"XXXX" + blob[4:])
    with pytest.raises(fhe.B200FheError):
        fhe.load_galois_keys(ctx, blob[:len(blob) // 2])
    other = fhe.Context(LOGN + 1, fhe.coeff_modulus_create(2 * N, BITS), NP)
    with pytest.raises(fhe.B200FheError):
        fhe.load_eval_key(other, fhe.save_eval_key(ctx, keys[2 * N - 1].digits))
    with pytest.raises(fhe.B200FheError):
        fhe.load_secret_key(other, fhe.save_secret_key(ctx, kg.secret()))
