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
