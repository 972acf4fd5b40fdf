"""GPU parity of the SEAL-style evaluator surface (negate, add/sub/multiply_plain, add_many,
multiply + relinearize, rescale_to_next, mod_switch_to_next, apply_galois, rotate with NAF decomposition, hoisting)
against the reference's own functions (src/evaluate.cu:95-2149) on the reference's keys (PhantomGaloisKey read
zero-copy): every limb bit-identical, scale and level equal.  Decrypted values: rotations and hoisting against numpy
rolls of the message, |err| < 2^-20 at scale 2^40."""
import ctypes as C

import numpy as np
import pytest

import refgpu

pytestmark = pytest.mark.gpu

LOGN, NP = 12, 2
BITS = [60, 40, 40, 40, 60, 60]      # |Q| = 4, |P| = 2: dnum = 2 (the reference needs |P| to divide |Q|)
SCALE = 2.0 ** 40
OPS = dict(negate=0, add_plain=1, sub_plain=2, multiply_plain=3, add=4, sub=5, multiply=6, relinearize=7, rescale=8,
           mod_switch=9, apply_galois=10, rotate=11, hoisting=12)


class Env:
    pass


@pytest.fixture(scope="module")
def env(fhe):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    if not refgpu.available():
        pytest.skip("oracle/_ref/libphantom_ref.so not built")
    e = Env()
    e.torch, e.fhe, e.L = torch, fhe, refgpu.lib()
    e.ref = refgpu.Ref(LOGN, BITS, NP)
    e.N, e.Q = 1 << LOGN, len(BITS) - NP
    e.ctx = fhe.Context(LOGN, e.ref.primes, NP)
    e.ev = fhe.Evaluator(e.ctx)
    L = e.L
    e.rlk = L.rref_gen_relinkey(e.ref.h)
    ptrs = (refgpu.vp * 64)()
    n = L.rref_relinkey_digits(e.ref.h, e.rlk, ptrs, 64)
    e.ev.set_relin_key([ptrs[d] for d in range(n)])
    e.gk = L.rref_gen_galois_keys(e.ref.h)
    assert e.gk, L.rref_last_error()
    elts = (C.c_uint32 * 256)()
    ne = L.rref_galois_elts(e.ref.h, elts, 256)
    e.elts = list(elts)[:ne]
    for i, g in enumerate(e.elts):
        n = L.rref_relinkey_digits(e.ref.h, L.rref_galois_key_get(e.gk, i), ptrs, 64)
        e.ev.set_seal_galois_key(g, [ptrs[d] for d in range(n)])
    e.kg = fhe.KeyGenerator(e.ctx, seed=bytes(range(32)))
    sk = torch.zeros(e.ref.nQP * e.N, dtype=torch.int64, device="cuda")
    refgpu.ck(L.rref_secret_key(e.ref.h, sk.data_ptr()))
    e.kg.set_secret(sk)
    e.enc = fhe.Encoder(e.ctx)
    return e


def fresh(e, seed, chain=1):
    x = np.random.default_rng(seed).uniform(-1, 1, e.N // 2)
    rct = e.L.rref_encrypt(e.ref.h, x.ctypes.data_as(C.POINTER(C.c_double)), len(x), SCALE, chain)
    assert rct, e.L.rref_last_error()
    return x, rct


def ours(e, rct):
    L = e.L
    return e.ev.import_ct(e.ref.ct_tensor(rct), L.rref_ct_limbs(rct), polys=L.rref_ct_size(rct), scale=L.rref_ct_scale(rct),
                          noise=L.rref_ct_noise(rct))


def same(e, o, rct, what):
    L = e.L
    e.ref.sync()
    e.torch.cuda.synchronize()
    d = o.desc()
    assert d.limbs == L.rref_ct_limbs(rct), "%s: limbs %d vs %d" % (what, d.limbs, L.rref_ct_limbs(rct))
    assert d.polys == L.rref_ct_size(rct), what + ": polynomial count"
    assert d.scale == L.rref_ct_scale(rct), "%s: scale %r vs %r" % (what, d.scale, L.rref_ct_scale(rct))
    a, b = o.tensor(), e.ref.ct_tensor(rct)
    e.torch.cuda.synchronize()
    if not bool((a.reshape(-1) == b.reshape(-1)).all()):
        bad = (a != b).reshape(d.polys, d.limbs, -1).any(dim=2).nonzero().tolist()
        raise AssertionError("%s: limbs differ from the reference at (poly, limb) %s" % (what, bad[:8]))


def rop(e, name, a, b=None, keys=None, arg=0, steps=()):
    arr = (C.c_int * max(len(steps), 1))(*steps)
    refgpu.ck(e.L.rref_seal_op(e.ref.h, OPS[name], a, b, keys, arg, arr, len(steps)))


def decrypt(e, o):
    d = o.desc()
    pt = e.kg.decrypt(o.tensor(), d.size_Ql, polys=d.polys)
    return e.enc.decode(pt, d.size_Ql, d.scale).real


def plain(e, seed, chain, scale=SCALE):
    v = np.random.default_rng(seed).uniform(-1, 1, e.N // 2)
    p = e.L.rref_encode_plain(e.ref.h, v.ctypes.data_as(C.POINTER(C.c_double)), len(v), scale, chain)
    assert p, e.L.rref_last_error()
    limbs = e.Q - (chain - 1)
    t = refgpu.as_tensor(e.L.rref_plain_data(p), limbs * e.N).clone()
    return v, p, t


def test_plain_ops_negate_add_many(env):
    e = env
    x, rct = fresh(e, 1)
    o = ours(e, rct)
    v, p, t = plain(e, 2, 1)
    for name, sub in (("add_plain", False), ("sub_plain", True)):
        rop(e, name, rct, p)
        e.ev.add_plain(o, t, SCALE, sub=sub)
        same(e, o, rct, name)
    rop(e, "negate", rct)
    e.ev.negate(o)
    same(e, o, rct, "negate_inplace")
    rop(e, "multiply_plain", rct, p)
    e.ev.multiply_plain(o, t, SCALE)
    same(e, o, rct, "multiply_plain_inplace")
    assert np.abs(decrypt(e, o) - (-x) * v).max() < 2.0 ** -20
    with pytest.raises(e.fhe.B200FheError):   # scale mismatch, like the reference's invalid_argument
        e.ev.add_plain(o, t, SCALE)
    # add_many of four ciphertexts
    xs, rcts = zip(*[fresh(e, 10 + i) for i in range(4)])
    arr = (refgpu.vp * 4)(*rcts)
    rsum = e.L.rref_add_many(e.ref.h, arr, 4)
    assert rsum, e.L.rref_last_error()
    osum = e.ev.add_many([ours(e, r) for r in rcts])
    same(e, osum, rsum, "add_many")
    assert np.abs(decrypt(e, osum) - sum(xs)).max() < 2.0 ** -20


def test_multiply_relinearize_rescale_modswitch(env):
    e = env
    (x, ra), (y, rb) = fresh(e, 3), fresh(e, 4)
    oa, ob = ours(e, ra), ours(e, rb)
    rop(e, "multiply", ra, rb)
    o3 = e.ev.multiply(oa, ob)
    same(e, o3, ra, "multiply_inplace (3 polynomials)")
    rop(e, "relinearize", ra, keys=e.rlk)
    o2 = e.ev.relinearize(o3)
    same(e, o2, ra, "relinearize_inplace")
    rop(e, "rescale", ra)
    o1 = e.ev.rescale_to_next(o2)
    same(e, o1, ra, "rescale_to_next")
    assert np.abs(decrypt(e, o1) - x * y).max() < 2.0 ** -20
    rop(e, "mod_switch", ra)
    o0 = e.ev.mod_switch_to_next(o1)
    same(e, o0, ra, "mod_switch_to_next")
    assert np.abs(decrypt(e, o0) - x * y).max() < 2.0 ** -20
    # squaring path (the same handle on both sides)
    x2, rc = fresh(e, 5)
    oc = ours(e, rc)
    rop(e, "multiply", rc, rc)
    same(e, e.ev.multiply(oc, oc), rc, "multiply_inplace (square)")


@pytest.mark.parametrize("chain", [1, 2])
def test_rotate_and_hoisting(env, chain):
    e = env
    x, rct = fresh(e, 20 + chain, chain)
    o = ours(e, rct)
    # a key of its own (power of two), NAF decompositions (3 = 4 - 1, 7 = 8 - 1, -5 = -4 - 1), and a long one
    for step in (1, -2, 3, 7, -5, 100):
        r2 = e.L.rref_ct_clone(rct)
        rop(e, "rotate", r2, keys=e.gk, arg=step)
        orot = e.ev.rotate(o, step)
        same(e, orot, r2, "rotate_inplace step %d" % step)
        assert np.abs(decrypt(e, orot) - np.roll(x, -step)).max() < 2.0 ** -20
    g = e.elts[3]
    r3 = e.L.rref_ct_clone(rct)
    rop(e, "apply_galois", r3, keys=e.gk, arg=g)
    same(e, e.ev.apply_galois(o, g), r3, "apply_galois_inplace elt %d" % g)
    steps = [1, 2, -4, 8, 16]
    r4 = e.L.rref_ct_clone(rct)
    rop(e, "hoisting", r4, keys=e.gk, steps=steps)
    oh = e.ev.hoisting(o, steps)
    same(e, oh, r4, "hoisting_inplace")
    assert np.abs(decrypt(e, oh) - sum(np.roll(x, -s) for s in steps)).max() < 2.0 ** -19


def test_our_seal_galois_keys_rotate_reference_ciphertexts(env):
    """keys from OUR generator (galois_key_seal) under the reference's secret: rotations decrypt correctly"""
    e = env
    fhe = e.fhe
    ev2 = fhe.Evaluator(e.ctx)
    keep = []
    for g in e.elts:
        k = e.kg.galois_key_seal(g)
        keep.append(k)
        ev2.set_seal_galois_key(g, k.digits)
    x, rct = fresh(e, 31, 2)
    L = e.L
    o = ev2.import_ct(e.ref.ct_tensor(rct), L.rref_ct_limbs(rct), scale=L.rref_ct_scale(rct))
    for step in (1, 5, -9):
        d = ev2.rotate(o, step)
        dd = d.desc()
        pt = e.kg.decrypt(d.tensor(), dd.size_Ql)
        assert np.abs(e.enc.decode(pt, dd.size_Ql, dd.scale).real - np.roll(x, -step)).max() < 2.0 ** -20
    h = ev2.hoisting(o, [4, -1])
    dd = h.desc()
    pt = e.kg.decrypt(h.tensor(), dd.size_Ql)
    assert np.abs(e.enc.decode(pt, dd.size_Ql, dd.scale).real - (np.roll(x, -4) + np.roll(x, 1))).max() < 2.0 ** -19
