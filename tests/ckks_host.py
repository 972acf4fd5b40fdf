"""Host-side CKKS decode used by the tests to state decrypted-value tolerances (TEST INFRASTRUCTURE).

The reference's own decoder cannot be used as the checker on this toolchain: its CRT compose
(src/rns_base.cu:172-236) adds multi-word integers with a carry chain split over separate asm statements
(include/uintmath.cuh:326-336), which nvcc 12.9 -O3 breaks for more than two limbs (measured on the B200:
encode->decode of a fresh plaintext returns ~1e83 for 6 limbs, 5e-11 for 2 limbs).  So the decrypted plaintext limbs
are taken from the reference's decrypt (c0 + c1*s, all on the hot-path arithmetic that IS bit-exact) and decoded
here: INTT (oracle), exact CRT with Python integers, centred lift, evaluation at zeta^(5^j) (the slot order of the
reference encoder, src/ckks.cu:103-180, src/fft.cu).
"""
import numpy as np


def crt_compose_centered(coeff_limbs, primes):
    """coeff_limbs [L][N] uint64 residues (coefficient form) -> list of N centred Python ints"""
    L, N = coeff_limbs.shape
    primes = [int(p) for p in primes[:L]]
    Q = 1
    for p in primes:
        Q *= p
    acc = [0] * N
    for i, p in enumerate(primes):
        Qi = Q // p
        w = Qi * pow(Qi % p, -1, p)
        row = coeff_limbs[i].tolist()
        for k in range(N):
            acc[k] += row[k] * w
    half = Q // 2
    out = []
    for v in acc:
        v %= Q
        out.append(v - Q if v > half else v)
    return out


def decode(o, pt_ntt_limbs, scale, nslots=None):
    """o: orc.Ctx; pt_ntt_limbs [L][N] in NTT form over primes 0..L-1 -> complex slots [N/2]"""
    L, N = pt_ntt_limbs.shape
    coeff = o.ntt_inv(np.ascontiguousarray(pt_ntt_limbs))
    ints = crt_compose_centered(coeff, o.primes)
    c = np.array([float(v) / scale for v in ints], dtype=np.float64)
    # m(zeta^t), zeta = exp(2 pi i / 2N): V[t] = sum_k c_k exp(+2 pi i t k / 2N) = conj(fft(c, 2N))[t] for real c
    V = np.conj(np.fft.fft(c, 2 * N))
    ns = N // 2 if nslots is None else nslots
    t, idx = 1, np.zeros(ns, dtype=np.int64)
    for j in range(ns):
        idx[j] = t
        t = t * 5 % (2 * N)
    return V[idx]
