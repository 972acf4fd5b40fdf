"""CPU: the oracle checked against independent mathematical definitions (SURVEY.md §8c plan (i))."""
import numpy as np
import pytest

import orc


@pytest.fixture(scope="module")
def small():
    primes = orc.coeff_modulus_create(1024, [60, 59, 59, 60, 60])
    return orc.Ctx(10, primes, 2)


def test_ntt_matches_quadratic_definition(small):
    rng = np.random.default_rng(7)
    x = orc.rand_residues(rng, small.primes, small.N)
    y = small.ntt_fwd(x)
    for i, q in enumerate(small.primes):
        psi = int(orc.lib().orc_minimal_primitive_root(2 * small.N, int(q)))
        assert (orc.ntt_fwd_naive(np.ascontiguousarray(x[i]), 10, int(q), psi) == y[i]).all()
    assert (small.ntt_inv(y) == x).all()


def test_ntt_pointwise_is_negacyclic_convolution(small):
    rng = np.random.default_rng(8)
    a = orc.rand_residues(rng, small.primes, small.N)
    b = orc.rand_residues(rng, small.primes, small.N)
    prod = small.binop("mul", small.size_QP, small.ntt_fwd(a), small.ntt_fwd(b))
    c = small.ntt_inv(prod)
    for i, q in enumerate(small.primes):
        assert (orc.negacyclic_mul_naive(np.ascontiguousarray(a[i]), np.ascontiguousarray(b[i]), int(q)) == c[i]).all()


def _crt(residues, primes):
    M = 1
    for p in primes:
        M *= int(p)
    x = 0
    for r, p in zip(residues, primes):
        Mi = M // int(p)
        x += int(r) * Mi * pow(Mi, -1, int(p))
    return x % M, M


def test_bconv_is_behz_approximate(small):
    """BEHZ fast conversion returns x + e*Q (mod p_j) with 0 <= e < |ibase| (no correction term)."""
    rng = np.random.default_rng(9)
    ib, ob = small.primes[:3], small.primes[3:]
    x = orc.rand_residues(rng, ib, 8)
    y = orc.bconv(x, ib, ob)
    for k in range(8):
        v, Q = _crt(x[:, k], ib)
        es = set()
        for j, p in enumerate(ob):
            found = [e for e in range(len(ib)) if (v + e * Q) % int(p) == int(y[j, k])]
            assert found, "not of the form x + e*Q"
            es.add(found[0])
        assert len(es) == 1


def test_keyswitch_decrypts(small):
    """A key-switching key for s' -> s built in numpy: ks(c2) must decrypt to c2*s' + small noise."""
    c = small
    rng = np.random.default_rng(10)
    N, Q, P, QP = c.N, c.size_Q, c.size_P, c.size_QP
    primes = [int(p) for p in c.primes]

    def small_poly(lo, hi):
        t = rng.integers(lo, hi, size=N)
        return np.stack([(t % p).astype(np.uint64) for p in primes])

    s, s2 = c.ntt_fwd(small_poly(-1, 2)), c.ntt_fwd(small_poly(-1, 2))
    beta = (Q + P - 1) // P
    Pmod = [1] * QP
    for i in range(QP):
        for p in primes[Q:]:
            Pmod[i] = Pmod[i] * p % primes[i]
    Ps2 = c.binop("mul", QP, np.stack([np.full(N, Pmod[i], dtype=np.uint64) for i in range(QP)]), s2)
    evk = []
    for d in range(beta):
        a = orc.rand_residues(rng, c.primes, N)
        e = c.ntt_fwd(small_poly(-3, 4))
        b = c.binop("sub", QP, e, c.binop("mul", QP, a, s))           # b = -a*s + e
        for i in range(d * P, min((d + 1) * P, Q)):                    # + P*s' on the digit's own limbs
            b[i] = c.binop("add", QP, b, Ps2)[i]
        evk.append(np.stack([b, a]))
    c2 = orc.rand_residues(rng, c.primes[:Q], N)
    ct = c.keyswitch(Q, np.zeros((2, Q, N), dtype=np.uint64), c2, evk)
    got = c.binop("add", Q, ct[0], c.binop("mul", Q, ct[1], s[:Q]))   # c0 + c1*s
    want = c.binop("mul", Q, c2, s2[:Q])
    diff = c.ntt_inv(c.binop("sub", Q, got, want))
    for i in range(Q):
        q = primes[i]
        cen = [int(x) - q if int(x) > q // 2 else int(x) for x in diff[i]]
        assert max(abs(x) for x in cen) < 2 ** 16, "key-switch noise too large: not a valid key switch"


def test_transform_of_a_polynomial_in_x_pow_step_repeats(small):
    """The identity behind ntt_forward_sub (csrc/ntt.cuh) and the compact weight plaintexts of DNN::conv: for
    p(X) = y(X^step), step = N / n', the first log2(n') stages of the N-point transform are the n'-point transform of
    y with the FIRST n' entries of the same twiddle table, and the remaining stages copy: NTT_N(p)[step * k + r] =
    NTT'_{n'}(y)[k] for every r < step.  Checked against the oracle's full transform, which is pinned to the reference."""
    rng = np.random.default_rng(11)
    N, logN = small.N, 10
    for sub_logn in (1, 4, 7, 10):
        n, step = 1 << sub_logn, N >> sub_logn
        y = orc.rand_residues(rng, small.primes, n)                 # [limbs][n']
        x = np.zeros((small.size_QP, N), dtype=np.uint64)
        x[:, ::step] = y
        full = small.ntt_fwd(x)
        for i, q in enumerate(small.primes):
            q = int(q)
            tw = [int(t) for t in orc.ntt_tables(logN, q)[0][:n]]   # psi powers in the reference's bit-reversed order
            a = [int(v) for v in y[i]]
            m, length = 1, n >> 1
            while m < n:                                            # Cooley-Tukey, natural in -> bit-reversed out
                for g in range(m):
                    w = tw[m + g]
                    for k in range(2 * g * length, 2 * g * length + length):
                        t = a[k + length] * w % q
                        a[k], a[k + length] = (a[k] + t) % q, (a[k] - t) % q
                m, length = 2 * m, length >> 1
            want = np.repeat(np.array(a, dtype=np.uint64), step)
            assert (full[i] == want).all(), "limb %d, n' = %d" % (i, n)


def test_lift_and_rescale_commute_with_the_transform(small):
    """What the fused plaintext path of DNN::conv relies on (csrc/encoder.cu decompose_lift_rescale_kernel): the per-limb
    scalar multiply and the rescale (a_j - [a_last]_{q_j}) * q_last^-1 are exact arithmetic mod q_j on each coefficient and
    the transform is linear, so doing both on the COEFFICIENTS and transforming Ql - 1 limbs gives the limbs of
    transform -> scalar multiply -> rescale (the oracle's restatement of the reference's divide_and_round_q_last_ntt)."""
    rng = np.random.default_rng(12)
    Ql, N = 3, small.N
    primes = [int(q) for q in small.primes[:Ql]]
    coeff = orc.rand_residues(rng, small.primes[:Ql], N)            # one polynomial, coefficient form, [Ql][N]
    lift = [int(rng.integers(1, q)) for q in primes]
    # the reference's order: NTT, multiply limb-wise by the scalar, rescale in NTT form
    ntt = small.ntt_fwd(coeff, 0, Ql)
    scaled = np.stack([np.array([int(v) * lift[j] % primes[j] for v in ntt[j]], dtype=np.uint64) for j in range(Ql)])
    want = small.rescale(Ql, scaled[None])[0]
    # coefficient domain first, one transform over Ql - 1 limbs afterwards
    inv = small.level_consts(Ql)["inv_q_last_mod_q"]
    last = [int(v) * lift[Ql - 1] % primes[Ql - 1] for v in coeff[Ql - 1]]
    rows = []
    for j in range(Ql - 1):
        q = primes[j]
        rows.append(np.array([((int(v) * lift[j] - t) % q) * int(inv[j]) % q for v, t in zip(coeff[j], last)], dtype=np.uint64))
    got = small.ntt_fwd(np.stack(rows), 0, Ql - 1)
    assert (got == want).all()


def test_partial_sums_over_ranks_are_exact(small):
    """What the multi-GPU modes rely on (DNN::conv over ranks, the baby steps of a limb-sharded stage): a 128-bit
    multiply-accumulate reduced once equals the modular sum of the ranks' separately reduced partial accumulations, whatever the
    split -- so dealing the summands to the ranks and adding the canonical partial results changes no bit.  Also the limb-wise
    claim behind the sharded linear-transform kernels: the oracle's inner product of one row depends on that row only."""
    rng = np.random.default_rng(21)
    q = int(small.primes[1])
    J = 37
    a = [int(v) for v in orc.rand_residues(rng, np.array([q] * J, dtype=np.uint64), 1).reshape(-1)]
    b = [int(v) for v in orc.rand_residues(rng, np.array([q] * J, dtype=np.uint64), 1).reshape(-1)]
    whole = sum(x * y for x, y in zip(a, b)) % q
    for G in (2, 3, 8):
        parts = [sum(x * y for x, y in zip(a[J * r // G:J * (r + 1) // G], b[J * r // G:J * (r + 1) // G])) % q for r in range(G)]
        acc = 0
        for p in parts:
            acc = (acc + p) % q     # addmod of canonical residues, as peer_row_sum does
        assert acc == whole
    # one row of the key-switch inner product from the oracle = the same row computed with every other row zeroed
    Ql, N = 3, small.N
    beta = (Ql + small.size_P - 1) // small.size_P
    QlP = Ql + small.size_P
    pr = np.concatenate([small.primes[:Ql], small.primes[len(small.primes) - small.size_P:]])
    up = np.stack([orc.rand_residues(rng, pr, N) for _ in range(beta)])
    evk = [orc.rand_residues(rng, small.primes, N, lead=(2,)) for _ in range(beta)]
    full = small.inner_prod(Ql, up, evk)
    t = 1
    masked = np.zeros_like(up)
    masked[:, t] = up[:, t]
    assert (small.inner_prod(Ql, masked, evk)[:, t] == full[:, t]).all()
