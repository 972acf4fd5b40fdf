"""CPU: host-side logic of the bootstrap layer (no GPU): collapsed-FFT parameters, rotation-index set, stage levels,
FLEXIBLEAUTO scaling factors, automorphism indices — against the values the reference derives for its own
bootstrapping_example configuration (src/util.cu:733-817, src/bootstrap.cu:595-826) and against an
independent Python restatement; and the no-CPU-fallback rule for the compute entries."""
import numpy as np
import pytest


def chain(fhe, logN, levels_after, nP):
    depth = levels_after + 18
    return fhe.coeff_modulus_create(1 << logN, [60] + [59] * depth + [60] * nP)


@pytest.fixture(scope="module")
def boot16(fhe):
    primes = chain(fhe, 16, 11, 10)
    ctx = fhe.Context(16, primes, 10, host_only=True)
    return fhe, ctx, fhe.Bootstrapper(ctx, (2, 2))


def test_collapsed_fft_params_reference_config(boot16):
    _, _, bt = boot16
    # {levelBudget, layersCollapse, remCollapse, numRotations, b, g, numRotationsRem, bRem, gRem}
    assert bt.params(0) == [2, 8, 7, 511, 8, 64, 255, 8, 32]
    assert bt.params(1) == [2, 8, 7, 511, 8, 64, 255, 8, 32]


def test_stage_shapes_and_levels(boot16):
    _, ctx, bt = boot16
    Q = ctx.size_Q
    assert Q == 30
    # CoeffsToSlots: stage 1 runs first at chain index 2 (|Q|-1 primes), the remainder stage 0 one level below
    assert bt.stage_info(0, 1) == (2, 511, Q - 1)
    assert bt.stage_info(0, 0) == (2, 255, Q - 2)
    # SlotsToCoeffs: after EvalMod (depth 8 + 6 double-angle levels) and the C2S budget
    assert bt.stage_info(1, 0) == (2, 511, Q - 16)
    assert bt.stage_info(1, 1) == (2, 255, Q - 17)


def _reduce(index, slots):
    return index % slots


def test_rotation_indices_match_python_restatement(boot16):
    fhe, ctx, bt = boot16
    N = ctx.N
    M, slots = 2 * N, N // 2
    lb, lc, rc, nr, b, g, nrr, br, gr = bt.params(0)
    want = set()
    for s in range(lb - 1, 0, -1):  # encode stages above the remainder
        sh = 1 << ((s - 1) * lc + rc)
        want |= {_reduce((j - (nr + 1) // 2 + 1) * sh, slots) for j in range(g)}
        want |= {_reduce(g * i * sh, M // 4) for i in range(b)}
    want |= {_reduce(j - (nrr + 1) // 2 + 1, slots) for j in range(gr)} | {_reduce(gr * i, M // 4) for i in range(br)}
    for s in range(lb - 1):         # decode stages
        sh = 1 << (s * lc)
        want |= {_reduce((j - (nr + 1) // 2 + 1) * sh, M // 4) for j in range(g)}
        want |= {_reduce(g * i * sh, M // 4) for i in range(b)}
    sh = 1 << ((lb - 1) * lc)
    want |= {_reduce((j - (nrr + 1) // 2 + 1) * sh, M // 4) for j in range(gr)}
    want |= {_reduce(gr * i * sh, M // 4) for i in range(br)}
    want -= {0, M // 4}
    got = bt.rotation_indices()
    assert got == sorted(want)
    assert len(got) == 170   # + the conjugation key = 171 evaluation keys (profiles/r01_bootstrap_full_parity.json)


def test_auto_index(fhe):
    for logN in (12, 16):
        m = 2 << logN
        for r in (1, 2, 7, 100, 4095):
            assert fhe.auto_index(logN, r) == pow(5, r, m)
            assert fhe.auto_index(logN, -r) == pow(pow(5, -1, m), r, m)
        assert fhe.auto_index(logN, 0) == 1 and fhe.auto_index(logN, m - 1) == m - 1


def test_scaling_factors(boot16):
    _, ctx, bt = boot16
    sf = bt.ev.scaling_factors(0)
    q = [float(int(p)) for p in ctx.primes[:ctx.size_Q]]
    want = [q[-1]]
    for k in range(1, len(q)):
        want.append(want[-1] * want[-1] / q[len(q) - k])
    assert sf == want
    assert bt.ev.scaling_factors(1) == [v * v for v in want[:-1]]


def test_bootstrap_has_no_cpu_fallback(boot16):
    fhe, ctx, bt = boot16
    with pytest.raises(fhe.B200FheError, match="no CPU fallback"):
        bt.ev.import_ct(12345, 3)   # any compute entry on a host-only context


def test_level_budget_1_1_rejected(fhe):
    primes = chain(fhe, 12, 2, 7)
    ctx = fhe.Context(12, primes, 7, host_only=True)
    with pytest.raises(fhe.B200FheError):
        fhe.Bootstrapper(ctx, (1, 1))


def test_peer_shard_arguments_and_arena_size(boot16):
    """b200fhe_boot_peer_arena_bytes / b200fhe_boot_set_peer_shard on a host-only bootstrapper: the arena of the limb-sharded
    bootstrap = two halves (inner sums of the widest stage b x [2][QP][N], a slot of partial rows per rank, the summed rows) +
    the EvalMod swap block + flag rows; bad rank / arena arguments are refused; the virtual-rank form needs no arena."""
    fhe, ctx, bt = boot16
    N, QP, Q = ctx.N, ctx.size_Q + ctx.size_P, ctx.size_Q
    b = max(bt.params(0)[4], bt.params(0)[7], bt.params(1)[4], bt.params(1)[7])
    total = (2 * QP + Q) * N * 8
    half = b * 2 * QP * N * 8 + (total + 9 * N * 8) + total
    assert bt.peer_arena_bytes() == 2 * half + 2 * Q * N * 8 + 256
    with pytest.raises(fhe.B200FheError):
        bt.set_peer_shard(2, 2, [1, 1], bt.peer_arena_bytes(), 1)      # rank out of range
    with pytest.raises(fhe.B200FheError):
        bt.set_peer_shard(10, 0, [1] * 10, bt.peer_arena_bytes(), 1)   # more than 9 ranks
    with pytest.raises(fhe.B200FheError):
        bt.set_peer_shard(2, 0, [1, 1], 1024, 1)                       # arena too small
    with pytest.raises(fhe.B200FheError):
        bt.set_peer_shard(2, 0, [1, None], bt.peer_arena_bytes(), 1)   # a peer's arena is missing
    bt.set_peer_shard(3, virtual_ranks=True)                           # no arena needed
    bt.set_peer_shard(1)                                               # off


def test_sharding_ranges_cover_everything():
    """the runs the limb-sharded stage / the rank-sharded layers deal out: contiguous, disjoint, complete, also when the
    count does not divide (same formulas as Bootstrapper::bsgs_stage and DNN::run_on_ranks: [n r / G, n (r + 1) / G))"""
    for n in (1, 7, 8, 16, 39, 40, 110):
        for G in range(1, 10):
            runs = [(n * r // G, n * (r + 1) // G) for r in range(G)]
            assert runs[0][0] == 0 and runs[-1][1] == n
            assert all(a[1] == b_[0] for a, b_ in zip(runs, runs[1:]))
            assert max(hi - lo for lo, hi in runs) - min(hi - lo for lo, hi in runs) <= 1
