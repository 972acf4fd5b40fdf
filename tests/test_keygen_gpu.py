"""Key generation, encryption and decryption on the B200.

Keys are random, so parity with the reference is FUNCTIONAL here and is checked in both directions
(test_cross_with_reference): the reference decrypts what we encrypt under its secret key, its evaluator multiplies /
relinearises / rotates with evaluation keys WE generated, and our decryption of its ciphertexts is bit-identical to its
own (decryption is deterministic).  The structure of every key is checked exactly: c0 + c1 * s' - P * s_old on the
digit's limbs must be a polynomial with centred-binomial coefficients (|e| <= 21), the same signed value under every
prime.  Floating-point tolerances are stated where they are used.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SEED = bytes(range(32))


def _gpu():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    return torch


def _mk(fhe, logN, bits, nP, seed=SEED):
    primes = fhe.coeff_modulus_create(1 << logN, bits)
    ctx = fhe.Context(logN, primes, nP)
    return ctx, fhe.KeyGenerator(ctx, seed)


def _centred(fhe, ctx, t, limbs, prime_idx=None):
    """coefficient form of an NTT-form polynomial [limbs][N] -> signed int64 per limb (host)"""
    x = t.clone()
    ctx.ntt_inverse(x, limbs)
    h = fhe.to_host(x).astype(np.uint64).reshape(limbs, ctx.N)
    out = np.empty((limbs, ctx.N), dtype=np.int64)
    for l in range(limbs):
        q = int(ctx.primes[l if prime_idx is None else prime_idx[l]])
        out[l] = np.where(h[l] > q // 2, h[l].astype(np.int64) - np.int64(q), h[l].astype(np.int64))
    return out


def test_chacha20_block_known_answers(fhe):
    _gpu()
    ctx, _ = _mk(fhe, 10, [40, 40], 1)
    # RFC 8439 section 2.3.2
    key = [int.from_bytes(bytes(range(4 * i, 4 * i + 4)), "little") for i in range(8)]
    got = fhe.chacha20_block(ctx, key, 1, [0x09000000, 0x4A000000, 0])
    want = [0xE4E7F110, 0x15593BD1, 0x1FDD0F50, 0xC47120A3, 0xC7F4D1C7, 0x0368C033, 0x9AAA2204, 0x4E6CD4C3,
            0x466482D2, 0x09AA9F07, 0x05D7C214, 0xA2028BD9, 0xD19C12B5, 0xB94E16DE, 0xE883D0CB, 0x4E3C50A2]
    assert got == want
    # random keys / counters / nonces against an independent implementation
    try:
        from cryptography.hazmat.primitives.ciphers import Cipher, algorithms
    except Exception:
        return
    rng = np.random.default_rng(1)
    for _ in range(8):
        kb = rng.integers(0, 256, 32, dtype=np.uint8).tobytes()
        nb = rng.integers(0, 256, 12, dtype=np.uint8).tobytes()
        ctr = int(rng.integers(0, 2 ** 32))
        ks = Cipher(algorithms.ChaCha20(kb, ctr.to_bytes(4, "little") + nb), mode=None).encryptor().update(bytes(64))
        got = fhe.chacha20_block(ctx, list(np.frombuffer(kb, dtype="<u4")), ctr, list(np.frombuffer(nb, dtype="<u4")))
        assert got == list(np.frombuffer(ks, dtype="<u4"))


def test_sampler_distributions(fhe):
    torch = _gpu()
    logN, bits, nP = 15, [60, 50, 45, 60], 1
    ctx, kg = _mk(fhe, logN, bits, nP)
    N, L = ctx.N, len(bits)
    primes = [int(p) for p in ctx.primes]

    def signed(t):
        h = fhe.to_host(t).astype(np.uint64).reshape(L, N)
        rows = []
        for l in range(L):
            q = primes[l]
            rows.append(np.where(h[l] > q // 2, h[l].astype(np.int64) - np.int64(q), h[l].astype(np.int64)))
        return np.stack(rows)

    # ternary (sample_ternary_poly, src/prng.cu:141-159): {-1, 0, 1}, the same value under every prime, p = 1/3 each
    t = signed(kg.sample(0, L))
    assert set(np.unique(t)) == {-1, 0, 1}
    assert all(np.array_equal(t[0], t[l]) for l in range(1, L))
    for v in (-1, 0, 1):   # 6 sigma of a binomial(N, 1/3)
        assert abs(np.count_nonzero(t[0] == v) - N / 3) < 6 * np.sqrt(N * 2 / 9)
    # centred binomial (sample_error_poly, :226-241): 21 + 21 coins, variance 10.5
    e = signed(kg.sample(1, L))
    assert np.abs(e).max() <= 21 and all(np.array_equal(e[0], e[l]) for l in range(1, L))
    assert abs(e[0].mean()) < 6 * np.sqrt(10.5 / N)
    assert abs(e[0].var() - 10.5) < 0.5
    # uniform (sample_uniform_poly, :171-205): canonical residues, flat histogram (chi-square over 64 cells,
    # 63 degrees of freedom: mean 63, sigma 11.2 -> < 140), limbs independent
    u = fhe.to_host(kg.sample(2, L)).astype(np.uint64).reshape(L, N)
    for l in range(L):
        assert int(u[l].max()) < primes[l]
        cells = np.floor(u[l].astype(np.float64) / primes[l] * 64).astype(np.int64).clip(0, 63)
        chi = ((np.bincount(cells, minlength=64) - N / 64) ** 2 / (N / 64)).sum()
        assert chi < 140, (l, chi)
    assert abs(np.corrcoef(u[0].astype(np.float64), u[1].astype(np.float64))[0, 1]) < 6 / np.sqrt(N)
    # streams: the next polynomial differs; the same seed reproduces the whole sequence; another seed does not
    u2 = fhe.to_host(kg.sample(2, L)).reshape(L, N)
    assert not np.array_equal(u2.view(np.uint64), u)
    kg_b = fhe.KeyGenerator(ctx, SEED)
    assert torch.equal(kg_b.secret(), kg.secret())
    assert np.array_equal(signed(kg_b.sample(0, L)), t)
    kg_c = fhe.KeyGenerator(ctx, bytes(range(1, 33)))
    assert not torch.equal(kg_c.secret(), kg.secret())
    kg_d, kg_e = fhe.KeyGenerator(ctx), fhe.KeyGenerator(ctx)     # operating-system entropy
    assert not torch.equal(kg_d.secret(), kg_e.secret())
    # the secret key itself: ternary in coefficient form
    s = _centred(fhe, ctx, kg.secret(), L)
    assert set(np.unique(s)) == {-1, 0, 1} and all(np.array_equal(s[0], s[l]) for l in range(1, L))
    # the Q_l P prime map: limbs 0..Ql-1 then the special primes
    e2 = fhe.to_host(kg.sample(1, 3, 2)).astype(np.uint64).reshape(3, N)
    neg = e2[2] > 2 ** 40
    assert neg.any() and np.array_equal(e2[2][neg] - np.uint64(primes[3]), e2[0][neg] - np.uint64(primes[0]))


@pytest.mark.parametrize("logN,bits,nP", [(12, [60, 50, 50, 50, 60, 60], 2), (13, [55, 45, 45, 45, 45, 55, 55, 55], 3),
                                          (12, [60, 50, 50, 60], 1)])
def test_key_structure(fhe, logN, bits, nP):
    """every digit of every key: c0 + c1 * s_new - [limb in digit] P * s_old = -e with e centred binomial"""
    torch = _gpu()
    import refgpu     # (only its raw-pointer -> tensor view; the reference library is not loaded here)
    ctx, kg = _mk(fhe, logN, bits, nP)
    N, QP, Q = ctx.N, len(bits), len(bits) - nP
    primes = [int(p) for p in ctx.primes]
    P = 1
    for p in primes[Q:]:
        P *= p
    s = kg.secret()
    s2 = torch.empty_like(s)
    ctx.eltwise(2, s2, s, s, QP)
    dnum = -(-Q // nP)

    def check(key, s_old, s_new):
        assert len(key.digits) == dnum
        for d, ptr in enumerate(key.digits):
            dig = refgpu.as_tensor(ptr, 2 * QP * N).view(2, QP, N)
            acc = torch.empty_like(s)
            ctx.eltwise(2, acc, dig[1].contiguous(), s_new, QP)
            ctx.eltwise(0, acc, acc, dig[0].contiguous(), QP)
            inj = torch.zeros_like(s)
            for l in range(d * nP, min((d + 1) * nP, Q)):
                tmp = torch.zeros_like(s)       # the constant P mod q_l on limb l, times s_old
                tmp[l] = P % primes[l]
                prod = torch.zeros_like(s)
                ctx.eltwise(2, prod, tmp, s_old, QP)
                inj[l] = prod[l]
            ctx.eltwise(1, acc, acc, inj, QP)
            e = _centred(fhe, ctx, acc, QP)
            assert np.abs(e).max() <= 21, "digit %d: noise is not a centred-binomial polynomial" % d
            assert all(np.array_equal(e[0], e[l]) for l in range(1, QP))
            assert 5.0 < e[0].var() < 16.0

    check(kg.relin_key(), s2, s)
    for rot in (1, -3, N // 4):
        k = fhe.auto_index(logN, rot)
        kinv = pow(k, -1, 2 * N)
        perm = torch.empty(N, dtype=torch.int32, device="cuda")
        ctx.automap_table(kinv, perm)
        s_rot = torch.empty_like(s)
        ctx.apply_galois_ntt(s_rot, s, perm, QP)
        check(kg.galois_key(k), s, s_rot)
    # public key: an encryption of zero over QP
    pk = kg.public_key()
    acc = torch.empty_like(s)
    ctx.eltwise(2, acc, pk[1].contiguous(), s, QP)
    ctx.eltwise(0, acc, acc, pk[0].contiguous(), QP)
    e = _centred(fhe, ctx, acc, QP)
    assert np.abs(e).max() <= 21 and e.any()


def test_encrypt_evaluate_decrypt_own_keys(fhe):
    """encode -> encrypt (symmetric and public-key) -> HMult + relinearise + rescale / rotate / conjugate with keys
    generated here -> decrypt -> decode.  Tolerance: 2^-25 at scale 2^40 (fresh noise ~ 2^-32 per slot)."""
    torch = _gpu()
    logN, bits, nP = 13, [60, 40, 40, 40, 40, 60, 60], 2
    ctx, kg = _mk(fhe, logN, bits, nP)
    enc = fhe.Encoder(ctx)
    N, Q = ctx.N, len(bits) - nP
    rng = np.random.default_rng(4)
    x = rng.uniform(-1, 1, N // 2) + 1j * rng.uniform(-1, 1, N // 2)
    y = rng.uniform(-1, 1, N // 2) + 1j * rng.uniform(-1, 1, N // 2)
    scale = 2.0 ** 40
    for Ql in (Q, Q - 2, 1):
        px = enc.encode(x, Ql, scale if Ql > 1 else 2.0 ** 30)
        ct = kg.encrypt_symmetric(px, Ql)
        got = enc.decode(kg.decrypt(ct, Ql), Ql, scale if Ql > 1 else 2.0 ** 30)
        assert np.abs(got - x).max() < (2.0 ** -25 if Ql > 1 else 2.0 ** -15), Ql
        # the ciphertext hides the plaintext: c0 - plain is not small
        assert not torch.equal(ct[0], px)
    pk = kg.public_key()
    for Ql in (Q, Q - 1):
        px = enc.encode(x, Ql, scale)
        ct = kg.encrypt_asymmetric(pk, px, Ql)
        got = enc.decode(kg.decrypt(ct, Ql), Ql, scale)
        assert np.abs(got - x).max() < 2.0 ** -22, Ql
    # evaluation
    cx = kg.encrypt_symmetric(enc.encode(x, Q, scale), Q)
    cy = kg.encrypt_symmetric(enc.encode(y, Q, scale), Q)
    rlk = kg.relin_key()
    prod = torch.zeros_like(cx)
    ctx.hmult_relin(Q, prod, cx, cy, rlk.digits)
    # the unrelinearised product decrypts with s^2
    t3 = torch.zeros((3, Q, N), dtype=torch.int64, device="cuda")
    ctx.tensor(Q, t3, cx, cy)
    got3 = enc.decode(kg.decrypt(t3, Q, polys=3), Q, scale * scale)
    assert np.abs(got3 - x * y).max() < 2.0 ** -22
    got = enc.decode(kg.decrypt(prod, Q), Q, scale * scale)
    assert np.abs(got - x * y).max() < 2.0 ** -22
    res = torch.zeros((2, Q - 1, N), dtype=torch.int64, device="cuda")
    ctx.rescale(Q, res, prod, 2)
    got = enc.decode(kg.decrypt(res, Q - 1), Q - 1, scale * scale / float(ctx.primes[Q - 1]))
    assert np.abs(got - x * y).max() < 2.0 ** -22
    keys = kg.rotation_keys([1, -7, 100])
    assert 2 * N - 1 in keys
    for r in (1, -7, 100):
        k = fhe.auto_index(logN, r)
        rot = torch.zeros_like(cx)
        ctx.rotate_fused(Q, rot, cx, k, keys[k].digits)
        got = enc.decode(kg.decrypt(rot, Q), Q, scale)
        assert np.abs(got - np.roll(x, -r)).max() < 2.0 ** -18, r     # key-switching noise at scale 2^40
    conj = torch.zeros_like(cx)
    ctx.rotate_fused(Q, conj, cx, 2 * N - 1, keys[2 * N - 1].digits)
    got = enc.decode(kg.decrypt(conj, Q), Q, scale)
    assert np.abs(got - np.conj(x)).max() < 2.0 ** -18
    # error behaviour
    with pytest.raises(fhe.B200FheError):
        kg.galois_key(4)
    with pytest.raises(fhe.B200FheError):
        kg.decrypt(cx, Q, polys=4)
    with pytest.raises(ValueError):
        fhe.KeyGenerator(ctx, b"short")


def test_cross_with_reference(fhe):
    """Both directions with the REFERENCE's secret key adopted by our generator."""
    torch = _gpu()
    import orc
    import refgpu
    if not refgpu.available():
        pytest.skip("oracle/_ref/libphantom_ref.so not built")
    logN, bits, nP = 13, [60, 50, 50, 50, 50, 50, 60, 60], 2
    ref = refgpu.Ref(logN, bits, nP)
    L = refgpu.lib()
    ctx = fhe.Context(logN, ref.primes, nP)
    kg = fhe.KeyGenerator(ctx, SEED)
    N, Q, QP = 1 << logN, len(bits) - nP, len(bits)
    sk = torch.empty((QP, N), dtype=torch.int64, device="cuda")
    refgpu.ck(L.rref_secret_key(ref.h, sk.data_ptr()))
    kg.set_secret(sk)
    torch.cuda.synchronize()
    o = orc.Ctx(logN, ref.primes, nP)
    enc = fhe.Encoder(ctx)
    rng = np.random.default_rng(31)
    va, vb = rng.uniform(-1, 1, N // 2), rng.uniform(-1, 1, N // 2)
    scale = 2.0 ** 50
    dp = lambda v: v.ctypes.data_as(C.POINTER(C.c_double))
    # (1) their ciphertext, our decryption: bit-identical to their decryption
    a = L.rref_encrypt(ref.h, dp(va), N // 2, scale, 1)
    assert a, L.rref_last_error()
    theirs = torch.zeros((Q, N), dtype=torch.int64, device="cuda")
    assert L.rref_decrypt_raw(ref.h, a, theirs.data_ptr()) == Q
    ours = kg.decrypt(ref.ct_tensor(a), Q)
    torch.cuda.synchronize()
    assert torch.equal(ours, theirs), "decryption differs from the reference"
    # (2) our ciphertexts (symmetric and public-key), their decryption
    pb = enc.encode(vb + 0j, Q, scale)
    for ct in (kg.encrypt_symmetric(pb, Q), kg.encrypt_asymmetric(kg.public_key(), pb, Q)):
        torch.cuda.synchronize()
        b = L.rref_ct_from_raw(ref.h, 1, 2, ct.data_ptr(), scale)
        assert b, L.rref_last_error()
        assert np.abs(ref.decrypt_slots(b, o).real - vb).max() < 2.0 ** -30
    # (3) their evaluator with OUR relinearisation and rotation keys
    rlk = kg.relin_key()
    torch.cuda.synchronize()
    arr = (C.c_void_p * len(rlk.digits))(*rlk.digits)
    rk = L.rref_relinkey_from_raw(arr, len(rlk.digits))
    assert rk, L.rref_last_error()
    a2 = L.rref_ct_clone(a)
    refgpu.ck(L.rref_multiply_and_relin(ref.h, a2, b, rk))
    refgpu.ck(L.rref_rescale(ref.h, a2))
    assert np.abs(ref.decrypt_slots(a2, o).real - va * vb).max() < 2.0 ** -20
    rots = [1, 5, -3]
    keys = kg.rotation_keys(rots)
    torch.cuda.synchronize()
    gk = L.rref_rotkeys_new()
    for k, key in keys.items():
        arr = (C.c_void_p * len(key.digits))(*key.digits)
        refgpu.ck(L.rref_rotkeys_set_raw(gk, k, arr, len(key.digits)))
    for r in rots:
        outc = L.rref_ct_clone(a)
        refgpu.ck(L.rref_rotate_fused(ref.h, gk, a, outc, r))
        assert np.abs(ref.decrypt_slots(outc, o).real - np.roll(va, -r)).max() < 2.0 ** -20, r
        L.rref_ct_free(outc)


def test_bootstrap_self_contained(fhe):
    """The whole pipeline without the reference: our keys, our encoder, our precomputed diagonals, our encryption,
    EvalBootstrap, our decryption.  N = 2^12, level budget {2,2}; tolerance 2^-8 max / 12 bits mean precision (measured: 8.4e-5 max, 18.5 bits; the
    uniform-ternary secret at this toy ring size has a heavy-tailed approximation error)."""
    torch = _gpu()
    logN, nP, levels_after = 12, 7, 2
    bits = [60] + [59] * (levels_after + 18) + [60] * nP
    ctx, kg = _mk(fhe, logN, bits, nP, bytes(range(7, 39)))
    N, Q = ctx.N, len(bits) - nP
    enc = fhe.Encoder(ctx)
    boot = fhe.Bootstrapper(ctx, (2, 2))
    boot.precompute(enc)
    ev = boot.ev
    rlk = kg.relin_key()
    ev.set_relin_key(rlk.digits)
    keys = kg.rotation_keys(boot.rotation_indices())
    for k, key in keys.items():
        ev.set_galois_key(k, key.digits)
    rng = np.random.default_rng(12)
    x = rng.uniform(-1, 1, N // 2)
    scale = 2.0 ** 59
    ct = ev.import_ct(kg.encrypt_symmetric(enc.encode(x + 0j, Q, scale), Q), Q, scale=scale, noise=1)
    while ct.desc().size_Ql > 3 or ct.desc().noise_scale_deg != 2:
        ev.mult_const(ct, 1.0)
    out = boot.eval(ct)
    d = out.desc()
    assert d.size_Ql >= 1 + levels_after and d.size_Ql > 3
    got = enc.decode(kg.decrypt(out.tensor(), d.size_Ql), d.size_Ql, d.scale)
    err = np.abs(got.real - x)
    prec = -np.log2(err.mean() + 1e-300)
    print("self-contained bootstrap N=2^%d: levels 3 -> %d, max abs err %.3e, mean precision %.1f bits" % (
        logN, d.size_Ql, err.max(), prec))
    assert err.max() < 2.0 ** -8 and prec > 12
    del keys, rlk


def test_same_seed_gives_the_reference_keys_bit_for_bit(fhe):
    """Same-seed key parity (north_star "same keys and PRNG seed").  The reference draws every 64-byte
    PRNG seed from std::random_device; oracle/ref_shim.cu makes that source seedable (rref_seed_next_secret_key), and
    b200fhe_keygen_create_reference_prng consumes the same byte stream in the reference's order with the reference's
    samplers restated (src/prng.cu).  Secret key, public key, relinearisation key and the rotation / conjugation keys of
    EvalAtIndexKeyGen must then be identical words."""
    torch = _gpu()
    import refgpu
    if not refgpu.available():
        pytest.skip("oracle/_ref/libphantom_ref.so not built")
    L = refgpu.lib()
    if not hasattr(L, "rref_seed_next_secret_key"):
        pytest.skip("reference shim without the seedable random_device")
    L.rref_seed_next_secret_key.argtypes = [C.c_uint64]
    L.rref_gen_publickey.argtypes = [C.c_void_p, C.c_void_p]
    for logN, bits, nP, master in ((12, [60, 59, 59, 59, 60, 60], 2, 0xB200F4E0), (13, [60, 50, 50, 50, 50, 50, 60, 60, 60], 3, 7)):
        N, QP = 1 << logN, len(bits)
        try:
            L.rref_seed_next_secret_key(master)                   # the context's own draws (primitive roots) stay unseeded
            ref = refgpu.Ref(logN, bits, nP)                      # PhantomSecretKey: gen_secretkey
            sk_ref = torch.empty((QP, N), dtype=torch.int64, device="cuda")
            refgpu.ck(L.rref_secret_key(ref.h, sk_ref.data_ptr()))
            pk_ref = torch.empty((2, QP, N), dtype=torch.int64, device="cuda")
            refgpu.ck(L.rref_gen_publickey(ref.h, pk_ref.data_ptr()))
            rlk = L.rref_gen_relinkey(ref.h)
            rlk_ref = ref.relinkey_digits(rlk)
            rots = [1, -3, 5]
            idx = np.array(rots, dtype=np.int32)
            keys = L.rref_gen_rotkeys(ref.h, idx.ctypes.data_as(C.POINTER(C.c_int32)), len(rots))
            assert rlk and keys, L.rref_last_error()
            autos = [L.rref_auto_index(ref.h, r) for r in rots] + [2 * N - 1]
            gk_ref = [ref.relinkey_digits(L.rref_rotkey_get(keys, k)) for k in autos]
        finally:
            L.rref_set_random_seed_off()
        ctx = fhe.Context(logN, ref.primes, nP)
        kg = fhe.KeyGenerator(ctx, reference_prng_seed=master)
        assert torch.equal(kg.secret(), sk_ref), "secret key differs (N=2^%d)" % logN
        assert torch.equal(kg.public_key(), pk_ref), "public key differs"

        def digits(key):
            return [refgpu.as_tensor(p, 2 * QP * N).clone() for p in key.digits]
        ours = digits(kg.relin_key())
        assert len(ours) == len(rlk_ref) and all(torch.equal(a, b) for a, b in zip(ours, rlk_ref)), "relinearisation key differs"
        for k, want in zip(autos, gk_ref):
            got = digits(kg.galois_key(k))
            assert len(got) == len(want) and all(torch.equal(a, b) for a, b in zip(got, want)), "rotation key %d differs" % k
        kg.close()
        ctx.close()
