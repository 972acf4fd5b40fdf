"""GPU parity: every C-ABI entry point against the CPU oracle, bit-exact, on the same seeded inputs."""
import numpy as np
import pytest

import orc

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    return torch


def make(fhe, logN, bitsQ, nP, bitP=60):
    bits = list(bitsQ) + [bitP] * nP
    primes = orc.coeff_modulus_create(1 << logN, bits)
    return fhe.Context(logN, primes, nP), orc.Ctx(logN, primes, nP)


def eq(a, b):
    return np.array_equal(np.asarray(a).reshape(-1), np.asarray(b).reshape(-1))


@pytest.mark.parametrize("logN", [3, 8, 10, 11, 12, 13, 14, 15, 16, 17])   # < 12: the single-CTA form (fnwt_1d / inwt_1d);
# 17 (test/ntt_test.cu:139,149, ntt_bench.cu:117): one radix-2 stage around the two-pass kernels
def test_ntt_forward_inverse(fhe, logN):
    _torch()
    L = 3 if logN < 16 else 4
    ctx, o = make(fhe, logN, [60] + [59] * (L - 1), 0)
    N = 1 << logN
    rng = np.random.default_rng(100 + logN)
    x = orc.rand_residues(rng, o.primes, N)
    # edge values: 0, q-1
    x[0, :4] = [0, 1, o.primes[0] - 1, o.primes[0] - 2]
    d = fhe.to_dev(x)
    ctx.ntt_forward(d, L)
    y = fhe.to_host(d)
    assert eq(y, o.ntt_fwd(x)), "forward NTT differs from oracle"
    ctx.ntt_inverse(d, L)
    assert eq(fhe.to_host(d), x), "inverse(forward(x)) != x"
    # inverse of arbitrary data vs oracle
    d2 = fhe.to_dev(x)
    ctx.ntt_inverse(d2, L)
    assert eq(fhe.to_host(d2), o.ntt_inv(x))


def test_ntt_test_config_T(fhe):
    """BASELINE config 0: test/ntt_test — N=2^12, 3 limbs, forward then inverse, constant input (ntt_test.cu:97-121)
    with the reference's 50-bit primes and with 60-bit primes; batch 1 and 10 polys."""
    _torch()
    for bits in ([50, 50, 50], [60, 60, 60]):
        ctx, o = make(fhe, 12, bits, 0)
        for batch in (1, 10):
            x = np.full((batch, 3, 4096), 2, dtype=np.uint64)
            d = fhe.to_dev(x)
            ctx.ntt_forward(d, 3, npoly=batch, poly_stride=3 * 4096)
            y = fhe.to_host(d).reshape(batch, 3, 4096)
            assert eq(y[0], o.ntt_fwd(x[0])) and eq(y[-1], o.ntt_fwd(x[0]))
            ctx.ntt_inverse(d, 3, npoly=batch, poly_stride=3 * 4096)
            assert eq(fhe.to_host(d), x)


def test_ntt_1d_test_config(fhe):
    """test/ntt_test.cu:20-75 (fnwt_1d_opt / inwt_1d_opt): dim 2048, 50-bit primes, batches of polynomials, random
    input; forward against the oracle, inverse(forward) = identity"""
    _torch()
    ctx, o = make(fhe, 11, [50, 50], 0)
    rng = np.random.default_rng(7)
    x = np.stack([orc.rand_residues(rng, o.primes, 2048) for _ in range(5)])
    d = fhe.to_dev(x)
    ctx.ntt_forward(d, 2, npoly=5, poly_stride=2 * 2048)
    y = fhe.to_host(d).reshape(5, 2, 2048)
    for b in range(5):
        assert eq(y[b], o.ntt_fwd(x[b]))
    ctx.ntt_inverse(d, 2, npoly=5, poly_stride=2 * 2048)
    assert eq(fhe.to_host(d), x)


def test_ntt_limb_ranges_and_special_mod(fhe):
    """start/limbs sub-ranges and the QlP -> QP prime map (include_special_mod)."""
    _torch()
    ctx, o = make(fhe, 12, [60, 59, 59, 59, 59], 2)
    N, Ql = 4096, 3
    rng = np.random.default_rng(3)
    pr = np.concatenate([o.primes[:Ql], o.primes[5:]])
    x = orc.rand_residues(rng, pr, N)
    d = fhe.to_dev(x)
    ctx.ntt_forward(d, Ql + 2, 0, size_Ql=Ql)
    assert eq(fhe.to_host(d), o.ntt_fwd(x, 0, Ql))
    # only the P limbs, inverse
    d = fhe.to_dev(x)
    ctx.ntt_inverse(d, 2, Ql, size_Ql=Ql)
    ref = x.copy()
    ref[Ql:] = o.ntt_inv(x, Ql, Ql)[Ql:]
    assert eq(fhe.to_host(d), ref)
    # middle range, plain map
    x5 = orc.rand_residues(rng, o.primes[:5], N)
    d = fhe.to_dev(x5)
    ctx.ntt_forward(d, 2, 2)
    ref = x5.copy()
    ref[2:4] = o.ntt_fwd(x5[:4], 2)[2:4]
    assert eq(fhe.to_host(d), ref)
    # zero limbs is a no-op
    ctx.ntt_forward(d, 0, 0)
    with pytest.raises(fhe.B200FheError):
        ctx.ntt_forward(d, 9, 0)


def test_ntt_inverse_scale_and_fuse_moddown(fhe):
    _torch()
    ctx, o = make(fhe, 13, [60, 59, 59, 59], 2)
    N, Ql = 8192, 4
    rng = np.random.default_rng(4)
    k = o.level_consts(Ql)
    x = orc.rand_residues(rng, o.primes[:Ql], N)
    out = fhe.to_dev(np.zeros_like(x))
    ctx.ntt_inverse_scale(out, fhe.to_dev(x), Ql, 0, 0, fhe.to_dev(k["partQlHatInv"]), fhe.to_dev(k["partQlHatInv_shoup"]))
    ref = o.ntt_inv(x)
    for i in range(Ql):
        q = int(o.primes[i])
        ref[i] = np.array([int(v) * int(k["partQlHatInv"][i]) % q for v in ref[i]], dtype=np.uint64)
    assert eq(fhe.to_host(out), ref)
    # fuse_moddown: ct = (cx - NTT(delta)) * PInv
    cx = orc.rand_residues(rng, o.primes[:Ql], N)
    delta = orc.rand_residues(rng, o.primes[:Ql], N)
    ct = fhe.to_dev(np.zeros_like(cx))
    ctx.ntt_forward_fuse_moddown(ct, fhe.to_dev(cx), fhe.to_dev(k["bigPInv_mod_q"]), fhe.to_dev(k["bigPInv_mod_q_shoup"]),
                                 fhe.to_dev(delta), Ql)
    nd = o.ntt_fwd(delta)
    ref = np.zeros_like(cx)
    for i in range(Ql):
        q = int(o.primes[i])
        ref[i] = np.array([(int(a) - int(b)) % q * int(k["bigPInv_mod_q"][i]) % q for a, b in zip(cx[i], nd[i])], dtype=np.uint64)
    assert eq(fhe.to_host(ct), ref)


def rand_evk(rng, o, beta):
    return [orc.rand_residues(rng, o.primes, o.N, lead=(2,)) for _ in range(beta)]


KS_CASES = [
    # logN, Q bit sizes, nP, levels to test
    (12, [60, 59, 59, 59, 59, 59], 2, [6, 5, 3, 1]),      # alpha 2, uneven last digit at Ql=5,3,1
    (12, [60, 59, 59, 59], 1, [4, 2]),                    # alpha == 1 fast paths
    (12, [60, 50, 50, 50, 50, 50, 50], 3, [7, 4]),        # mixed prime sizes, digit of 1 limb
]


@pytest.mark.parametrize("case", KS_CASES)
def test_keyswitch_pipeline(fhe, case):
    _torch()
    logN, bitsQ, nP, levels = case
    ctx, o = make(fhe, logN, bitsQ, nP)
    N = 1 << logN
    rng = np.random.default_rng(11)
    for Ql in levels:
        beta = (Ql + nP - 1) // nP
        assert ctx.beta(Ql) == beta
        QlP = Ql + nP
        c2 = orc.rand_residues(rng, o.primes[:Ql], N)
        evk = rand_evk(rng, o, beta)
        devk = [fhe.to_dev(e) for e in evk]
        # modup
        up = fhe.to_dev(np.zeros((beta, QlP, N), dtype=np.uint64))
        ctx.modup(Ql, up, fhe.to_dev(c2))
        up_ref = o.modup(Ql, c2)
        assert eq(fhe.to_host(up), up_ref), "modup differs (Ql=%d)" % Ql
        # inner product
        cx = fhe.to_dev(np.zeros((2, QlP, N), dtype=np.uint64))
        ctx.inner_prod(Ql, cx, up, devk)
        cx_ref = o.inner_prod(Ql, up_ref, evk)
        assert eq(fhe.to_host(cx), cx_ref), "inner product differs (Ql=%d)" % Ql
        # moddown, out of place and in place (aliasing as eval_key_switch.cu:190)
        ct = fhe.to_dev(np.zeros((Ql, N), dtype=np.uint64))
        cx0 = fhe.to_dev(cx_ref[0])
        ctx.moddown_from_ntt(Ql, ct, cx0)
        md_ref = o.moddown(Ql, cx_ref[0])
        assert eq(fhe.to_host(ct), md_ref), "moddown differs (Ql=%d)" % Ql
        cx1 = fhe.to_dev(cx_ref[1])
        ctx.moddown_from_ntt(Ql, cx1, cx1)
        assert eq(fhe.to_host(cx1)[:Ql], o.moddown(Ql, cx_ref[1]))
        # whole key switch, accumulating into ct
        ct0 = orc.rand_residues(rng, o.primes[:Ql], N, lead=(2,))
        dct = fhe.to_dev(ct0)
        ctx.keyswitch(Ql, dct, fhe.to_dev(c2), devk)
        assert eq(fhe.to_host(dct), o.keyswitch(Ql, ct0, c2, evk)), "keyswitch differs (Ql=%d)" % Ql
        # P -> Ql conversion on its own
        if True:
            src = orc.rand_residues(rng, o.primes[o.size_Q:], N)
            dst = fhe.to_dev(np.zeros((Ql, N), dtype=np.uint64))
            ctx.bconv_P_to_Ql(Ql, dst, fhe.to_dev(src))
            if nP > 1:
                ref = orc.bconv(src, o.primes[o.size_Q:], o.primes[:Ql])
            else:
                p = int(o.primes[o.size_Q])
                ref = np.stack([np.where(p > int(q), src[0] % q, src[0]) for q in o.primes[:Ql]])
            assert eq(fhe.to_host(dst), ref)


def test_evaluator_ops(fhe):
    _torch()
    ctx, o = make(fhe, 12, [60, 59, 59, 59, 59], 2)
    N, L = 4096, 5
    rng = np.random.default_rng(12)
    a = orc.rand_residues(rng, o.primes[:L], N, lead=(2,))
    b = orc.rand_residues(rng, o.primes[:L], N, lead=(2,))
    da, db = fhe.to_dev(a), fhe.to_dev(b)
    out = fhe.to_dev(np.zeros((3, L, N), dtype=np.uint64))
    ctx.tensor(L, out, da, db)
    assert eq(fhe.to_host(out), o.tensor(L, a, b))
    ctx.tensor(L, out, da, da)
    assert eq(fhe.to_host(out), o.tensor_square(L, a))
    # element-wise
    for op, name in ((0, "add"), (1, "sub"), (2, "mul")):
        r = fhe.to_dev(np.zeros((L, N), dtype=np.uint64))
        ctx.eltwise(op, r, da, db, L)
        assert eq(fhe.to_host(r), o.binop(name, L, a[0], b[0]))
    r = fhe.to_dev(np.zeros((L, N), dtype=np.uint64))
    ctx.eltwise(3, r, da, None, L)
    assert eq(fhe.to_host(r), o.binop("sub", L, np.zeros((L, N), dtype=np.uint64), a[0]))
    # rescale, 2 and 3 polys, several levels
    for Lq in (5, 3, 2):
        for polys in (2, 3):
            src = orc.rand_residues(rng, o.primes[:Lq], N, lead=(polys,))
            dst = fhe.to_dev(np.zeros((polys, Lq - 1, N), dtype=np.uint64))
            dsrc = fhe.to_dev(src)
            ctx.rescale(Lq, dst, dsrc, polys)
            assert eq(fhe.to_host(dst), o.rescale(Lq, src)), "rescale differs (L=%d polys=%d)" % (Lq, polys)
            assert eq(fhe.to_host(dsrc), src), "rescale must not clobber its source"
    # automorphism tables and gather
    import torch
    for k in (5, 25, 2 * N - 1, pow(5, 77, 2 * N)):
        perm = torch.zeros(N, dtype=torch.int32, device="cuda")
        ctx.automap_table(k, perm)
        assert eq(perm.cpu().numpy().view(np.uint32), orc.automap_table(12, k))
        assert eq(fhe.galois_table(12, k), orc.galois_table(12, k))
    perm_h = orc.automap_table(12, 25)
    dst = fhe.to_dev(np.zeros_like(a))
    ctx.apply_galois_ntt(dst, da, torch.from_numpy(perm_h.view(np.int32)).cuda(), L, 2, L * N)
    assert eq(fhe.to_host(dst), a[:, :, perm_h])


@pytest.mark.parametrize("overlap,lanes", [(7, 1), (0, 1), (1, 1), (2, 1), (3, 1), (4, 1), (6, 1), (7, 0), (4, 0), (0, 0)])
def test_hmult_relin_and_rotate(fhe, overlap, lanes, monkeypatch):
    """overlap: the B200FHE_OVERLAP settings of the key switch (bit 0: (d0, d1) of the tensor product on a background stream,
    bit 1: Q_l rows of the inner product beside the moddown chains, bit 2: row pass of the modup transforms fused with the inner
    product); lanes 0: everything on the caller's stream.  Every setting must give the oracle's limbs."""
    _torch()
    monkeypatch.setenv("B200FHE_OVERLAP", str(overlap))
    ctx, o = make(fhe, 12, [60, 59, 59, 59, 59], 2)
    ctx.set_lanes(lanes)
    N = 4096
    rng = np.random.default_rng(13)
    for Ql in (5, 4, 2):
        beta = (Ql + 1) // 2
        a = orc.rand_residues(rng, o.primes[:Ql], N, lead=(2,))
        b = orc.rand_residues(rng, o.primes[:Ql], N, lead=(2,))
        evk = rand_evk(rng, o, beta)
        devk = [fhe.to_dev(e) for e in evk]
        out = fhe.to_dev(np.zeros_like(a))
        da, db = fhe.to_dev(a), fhe.to_dev(b)
        ctx.hmult_relin(Ql, out, da, db, devk)
        assert eq(fhe.to_host(out), o.hmult_relin(Ql, a, b, evk)), "HMult+relin differs (Ql=%d)" % Ql
        ctx.hmult_relin(Ql, out, da, da, devk)   # square path
        assert eq(fhe.to_host(out), o.hmult_relin(Ql, a, a, evk))
        ctx.hmult_relin(Ql, da, da, db, devk)    # in place
        assert eq(fhe.to_host(da), o.hmult_relin(Ql, a, b, evk))
        for k in (5, 2 * N - 1):
            ctx.rotate_fused(Ql, out, db, k, devk)
            assert eq(fhe.to_host(out), o.rotate_fused(Ql, b, evk, orc.automap_table(12, k))), "rotate differs"


def test_host_buffer_entry_points(fhe):
    _torch()
    ctx, o = make(fhe, 12, [60, 59, 59], 0)
    rng = np.random.default_rng(14)
    x = orc.rand_residues(rng, o.primes, 4096)
    h = x.copy()
    ctx.ntt_forward_host(h, 3)
    assert eq(h, o.ntt_fwd(x))
    ctx.ntt_inverse_host(h, 3)
    assert eq(h, x)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_limb_sharded_keyswitch_virtual_ranks(fhe, world):
    """limb-sharded key switch (SURVEY.md 8e mode 2) with every rank's share run on this GPU in turn: the union of the
    ranks' rows equals b200fhe_keyswitch bit for bit, for even and uneven limb partitions and at a lower level"""
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from phantom_fhe_boot_b200 import sharding
    logN, nP = 12, 4
    N = 1 << logN
    primes = fhe.coeff_modulus_create(N, [60] + [59] * 11 + [60] * nP)
    ctx = fhe.Context(logN, primes, nP)
    g = torch.Generator(device="cuda").manual_seed(5)

    def rnd(*shape):
        return torch.randint(0, 1 << 58, shape, dtype=torch.int64, device="cuda", generator=g)
    for Ql in (12, 7):
        beta = ctx.beta(Ql)
        evk = [rnd(2, 12 + nP, N) for _ in range(beta)]
        ct, c2 = rnd(2, Ql, N), rnd(Ql, N)
        want = ct.clone()
        ctx.keyswitch(Ql, want, c2, evk)
        ks = sharding.LimbShardedKeySwitch(ctx, Ql, world=world)
        got = ks.run(c2, evk, acc=ct, acc_polys=0)
        torch.cuda.synchronize()
        assert bool((got == want).all()), "limb-sharded key switch differs at |Q_l| = %d, %d ranks" % (Ql, world)
        # and the oracle (the CPU restatement of keyswitch_inplace) is the checker, not only the single-GPU product path
        o = orc.Ctx(logN, primes, nP)
        want_o = o.keyswitch(Ql, fhe.to_host(ct).reshape(2, Ql, N), fhe.to_host(c2).reshape(Ql, N),
                             [fhe.to_host(e).reshape(2, 12 + nP, N) for e in evk])
        assert np.array_equal(fhe.to_host(got).reshape(-1), np.asarray(want_o).reshape(-1)), \
            "limb-sharded key switch differs from the oracle at |Q_l| = %d, %d ranks" % (Ql, world)


@pytest.mark.gpu
def test_limb_sharded_keyswitch_peer_memory_two_gpus():
    """two processes, two GPUs: the limb-sharded key switch with its row exchanges done by b200fhe_peer_exchange over
    NVLink peer memory (no NCCL on the data path) equals the single-GPU key switch on every rank, also after
    back-to-back runs and as a replayed CUDA graph.  Skipped on a box with one GPU."""
    import json
    import os
    import subprocess
    import sys
    import torch
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    # store kernel / stores inside the transforms / store kernel through the NVSwitch multicast address (where available)
    for extra, name in (([], "peer-unfused"), (["--fused"], "peer-fused"), (["--multicast"], "peer-unfused")):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
               "127.0.0.1", "--master-port", "29631", os.path.join(root, "tools", "limb_shard_bench.py"), "--transport", "peer",
               "--graph", "--logn", "13", "--q", "12", "--p", "4", "--iters", "5"] + extra
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
        out = json.loads(line)
        assert out["bit_exact_vs_single_gpu_all_ranks"] is True and out["transport"].startswith(name)


def test_ntt_host_batch_pipeline(fhe):
    """b200fhe_ntt_host_batch (pinned host buffer in and out, limb runs through the copy / compute / copy pipeline): forward
    against the oracle for a limb count that does not divide into the runs evenly, and forward + inverse = identity"""
    import os
    os.environ.setdefault("B200FHE_HOST_CHUNKS", "4")   # read once per process, at the first call: ragged runs of 2, 2, 1 limbs
    torch = _torch()
    logN, L, batch = 12, 5, 3
    N = 1 << logN
    ctx, o = make(fhe, logN, [60, 59, 59, 59, 59], 0)
    rng = np.random.default_rng(31)
    x = np.stack([orc.rand_residues(rng, o.primes, N) for _ in range(batch)])
    h = torch.from_numpy(x.view(np.int64).copy()).pin_memory()
    ctx.ntt_host_batch(h.data_ptr(), L, batch, ops=1)
    got = h.numpy().view(np.uint64)
    for b in range(batch):
        assert eq(got[b], o.ntt_fwd(x[b])), "host-batch forward transform differs from the oracle (polynomial %d)" % b
    ctx.ntt_host_batch(h.data_ptr(), L, batch, ops=2)
    assert eq(h.numpy().view(np.uint64), x)
    ctx.ntt_host_batch(h.data_ptr(), L, batch, ops=3)
    assert eq(h.numpy().view(np.uint64), x)


def test_cluster_inverse_transform_bit_exact(fhe):
    """ntt_inv_cluster (one kernel, a cluster of 8 blocks per limb, transpose through distributed shared memory; selected with
    b200fhe_tune_cluster_ntt) against the oracle and against the two-pass kernels at N = 2^16, with and without the per-limb
    scale, in place and with a polynomial stride."""
    torch = _torch()
    logN, nP = 16, 2
    N = 1 << logN
    ctx, o = make(fhe, logN, [60, 59, 59], nP)
    L = fhe.lib()
    rng = np.random.default_rng(77)
    try:
        x = orc.rand_residues(rng, o.primes[:3], N, lead=(2,))
        want = np.stack([o.ntt_inv(x[z]) for z in range(2)])
        for knob in (0, 64):
            L.b200fhe_tune_cluster_ntt(knob)
            d = fhe.to_dev(x)
            ctx.ntt_inverse(d, 3, npoly=2, poly_stride=3 * N)
            assert eq(fhe.to_host(d).reshape(want.shape), want), "inverse transform differs from the oracle (knob %d)" % knob
        # the scaled inverse inside a key switch (digit chains) and the whole HMult+relin: cluster form == two-pass form
        Ql = 3
        a = orc.rand_residues(rng, o.primes[:Ql], N, lead=(2,))
        b = orc.rand_residues(rng, o.primes[:Ql], N, lead=(2,))
        rlk = [fhe.to_dev(orc.rand_residues(rng, o.primes, N, lead=(2,))) for _ in range(ctx.beta(Ql))]
        outs = []
        for knob in (0, 64):
            L.b200fhe_tune_cluster_ntt(knob)
            out = fhe.to_dev(np.zeros_like(a))
            ctx.hmult_relin(Ql, out, fhe.to_dev(a), fhe.to_dev(b), rlk)
            torch.cuda.synchronize()
            outs.append(fhe.to_host(out))
        assert eq(outs[0], outs[1])
    finally:
        L.b200fhe_tune_cluster_ntt(-1)


def test_shim_entries_tables_ranges_devkeys(fhe):
    """The entries the C++ shim (shim/) uses: a context on caller-owned device tables in the reference's DNTTTable
    layout, transforms over an explicit limb -> prime description (include_special_mod / exclude_range / out of place),
    and the key switch with the digit pointers in device memory; each against the oracle."""
    torch = _torch()
    logN, nP = 12, 2
    N = 1 << logN
    ctx0, o = make(fhe, logN, [60, 59, 59, 59, 59], nP)
    tabs = [fhe.to_dev(ctx0.table(w)) for w in range(6)]          # tw, tw_shoup, itw, itw_shoup, ninv, ninv_shoup
    ctx = fhe.Context.from_tables(logN, o.primes, 0, *tabs)
    with pytest.raises(fhe.B200FheError):
        ctx.table(0)                                               # no host copies of caller-owned tables
    rng = np.random.default_rng(21)
    size_QP, Ql = len(o.primes), 3
    QlP = Ql + nP
    pr = np.concatenate([o.primes[:Ql], o.primes[5:]])
    x = orc.rand_residues(rng, pr, N)
    # include_special_mod over the whole QlP buffer, in place
    d = fhe.to_dev(x)
    ctx.ntt_forward_ex(d, d, 0, QlP, special_from=Ql, special_base=size_QP - nP)
    want = o.ntt_fwd(x, 0, Ql)
    assert eq(fhe.to_host(d), want)
    # exclude_range [1, 2): that limb stays as it was
    d = fhe.to_dev(x)
    ctx.ntt_forward_ex(d, d, 0, QlP, special_from=Ql, special_base=size_QP - nP, excl=(1, 2))
    w2 = want.copy()
    w2[1] = x[1]
    assert eq(fhe.to_host(d), w2)
    # inverse of the P limbs only, out of place, with a per-limb scale
    k = o.level_consts(Ql)
    sc = np.zeros(QlP, dtype=np.uint64)
    scs = np.zeros(QlP, dtype=np.uint64)
    sc[Ql:], scs[Ql:] = k["PHatInv_mod_p"], k["PHatInv_mod_p_shoup"]
    out = fhe.to_dev(np.zeros_like(x))
    ctx.ntt_inverse_ex(out, fhe.to_dev(x), Ql, nP, special_from=Ql, special_base=size_QP - nP, scale=fhe.to_dev(sc),
                       scale_shoup=fhe.to_dev(scs))
    ref = o.ntt_inv(x, Ql, Ql)
    got = fhe.to_host(out).reshape(x.shape)
    for i in range(nP):
        q = int(o.primes[5 + i])
        assert eq(got[Ql + i], np.array([int(v) * int(sc[Ql + i]) % q for v in ref[Ql + i]], dtype=np.uint64))
    with pytest.raises(fhe.B200FheError):
        ctx.ntt_forward_ex(d, d, 0, QlP, special_from=Ql, special_base=size_QP)   # maps past the tables
    # the special-prime count arrives later (a DNTTTable does not know it); then the key switch, keys by device table
    with pytest.raises(fhe.B200FheError):
        ctx.modup(Ql, d, d)
    ctx.set_special(nP)
    beta = (Ql + nP - 1) // nP
    c2 = orc.rand_residues(rng, o.primes[:Ql], N)
    evk = rand_evk(rng, o, beta)
    devk = [fhe.to_dev(e) for e in evk]
    table = torch.tensor([e.data_ptr() for e in devk], dtype=torch.int64, device="cuda")
    up = fhe.to_dev(o.modup(Ql, c2))
    cx = fhe.to_dev(np.zeros((2, QlP, N), dtype=np.uint64))
    ctx.inner_prod_devkeys(Ql, cx, up, table)
    assert eq(fhe.to_host(cx), o.inner_prod(Ql, fhe.to_host(up).reshape(beta, QlP, N), evk))
    ct0 = orc.rand_residues(rng, o.primes[:Ql], N, lead=(2,))
    dct = fhe.to_dev(ct0)
    ctx.keyswitch_devkeys(Ql, dct, fhe.to_dev(c2), table)
    assert eq(fhe.to_host(dct), o.keyswitch(Ql, ct0, c2, evk))
