"""CPU: oracle/fhe_oracle.c and the product's host code (hostmath.cpp behind the C ABI) against golden vectors
generated from the REFERENCE's own host code (tests/golden/make_golden.py -> ref_host_golden.npz)."""
import hashlib
import os

import numpy as np
import pytest

import orc
from golden.make_golden import CONFIGS

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_host_golden.npz"))


def _sha(arrs):
    return np.frombuffer(b"".join(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest() for a in arrs), dtype=np.uint8)


def _picked(name, n):
    logN = CONFIGS[name][0]
    return [i for i in range(n) if not (logN == 16 and i not in (0, 1, n - 1))]


@pytest.mark.parametrize("name", list(CONFIGS))
def test_oracle_primes_and_tables(name):
    logN, bits, nP = CONFIGS[name]
    pr = orc.coeff_modulus_create(1 << logN, bits)
    assert (pr == G[name + "_primes"]).all()
    for j, i in enumerate(_picked(name, len(pr))):
        q = int(pr[i])
        tw, tws, itw, itws, ninv, ninvs = orc.ntt_tables(logN, q)
        assert int(orc.lib().orc_minimal_primitive_root(2 << logN, q)) == int(G[name + "_roots"][j])
        assert ninv == int(G[name + "_ninv"][j]) and ninvs == int(G[name + "_ninv_shoup"][j])
        assert (_sha([tw, tws, itw, itws]) == G[name + "_table_sha256"][j]).all()
    if name == "T60":
        tw, tws, itw, itws, _, _ = orc.ntt_tables(12, int(pr[0]))
        for k, a in zip(("tw", "tw_shoup", "itw", "itw_shoup"), (tw, tws, itw, itws)):
            assert (a == G["T60_q0_" + k]).all()


@pytest.mark.parametrize("name", [n for n in CONFIGS if CONFIGS[n][2]])
def test_oracle_level_constants(name):
    logN, bits, nP = CONFIGS[name]
    pr = G[name + "_primes"]
    # level constants do not depend on N: build the oracle context at a tiny degree with the same primes?
    # primes are NTT-friendly for 2^logN (hence for smaller degrees too), so logN=4 keeps this instant.
    ctx = orc.Ctx(4, pr, nP)
    Q = len(pr) - nP
    k = ctx.level_consts(Q)
    assert (k["partQlHatInv"][:nP] == G[name + "_d0_hatinv"]).all()
    assert (k["partQlHatInv_shoup"][:nP] == G[name + "_d0_hatinv_shoup"]).all()
    assert (k["modup_mats"][0] == G[name + "_d0_mat"]).all()
    assert (k["PHatInv_mod_p"] == G[name + "_P_hatinv"]).all()
    assert (k["PHatInv_mod_p_shoup"] == G[name + "_P_hatinv_shoup"]).all()
    assert (k["PHat_mod_q"] == G[name + "_P_mat"]).all()


@pytest.mark.parametrize("name", list(CONFIGS))
def test_product_host_tables(fhe, name):
    """hostmath.cpp (the product's own table generator) == reference host code, through the C ABI."""
    logN, bits, nP = CONFIGS[name]
    pr = fhe.coeff_modulus_create(1 << logN, bits)
    assert (pr == G[name + "_primes"]).all()
    picked = _picked(name, len(pr))
    sub = np.array([pr[i] for i in picked], dtype=np.uint64)
    ctx = fhe.Context(logN, sub, 0, host_only=True)
    tabs = [ctx.table(w) for w in range(9)]
    for j in range(len(picked)):
        assert int(tabs[4][j, 0]) == int(G[name + "_ninv"][j]) and int(tabs[5][j, 0]) == int(G[name + "_ninv_shoup"][j])
        assert (_sha([tabs[0][j], tabs[1][j], tabs[2][j], tabs[3][j]]) == G[name + "_table_sha256"][j]).all()
        assert int(tabs[7][j, 0]) == int(G[name + "_ratio"][j][0]) and int(tabs[8][j, 0]) == int(G[name + "_ratio"][j][1])
    ctx.close()


@pytest.mark.parametrize("name", [n for n in CONFIGS if CONFIGS[n][2]])
def test_product_level_constants(fhe, name):
    logN, bits, nP = CONFIGS[name]
    pr = G[name + "_primes"]
    ctx = fhe.Context(4, pr, nP, host_only=True)
    Q = len(pr) - nP
    assert (ctx.level_const(Q, 0)[:nP] == G[name + "_d0_hatinv"]).all()
    assert (ctx.level_const(Q, 1)[:nP] == G[name + "_d0_hatinv_shoup"]).all()
    assert (ctx.level_const(Q, 100).reshape(G[name + "_d0_mat"].shape) == G[name + "_d0_mat"]).all()
    assert (ctx.level_const(Q, 5) == G[name + "_P_hatinv"]).all()
    assert (ctx.level_const(Q, 6) == G[name + "_P_hatinv_shoup"]).all()
    assert (ctx.level_const(Q, 7).reshape(Q, nP) == G[name + "_P_mat"]).all()
    # every level of the chain agrees with the oracle (which is pinned above)
    octx = orc.Ctx(4, pr, nP)
    for Ql in (1, 2, nP, nP + 1, Q - 1, Q):
        if Ql < 1 or Ql > Q:
            continue
        k = octx.level_consts(Ql)
        assert ctx.beta(Ql) == k["beta"]
        assert (ctx.level_const(Ql, 0) == k["partQlHatInv"]).all()
        assert (ctx.level_const(Ql, 3) == k["bigPInv_mod_q"]).all()
        assert (ctx.level_const(Ql, 2) == k["bigP_mod_q"]).all()
        assert (ctx.level_const(Ql, 8) == k["inv_q_last_mod_q"]).all()
        for d in range(k["beta"]):
            assert (ctx.level_const(Ql, 100 + d) == k["modup_mats"][d].reshape(-1)).all()
    ctx.close()
